"""Port and scheduling constants the hot path reads (reference harness/determined/constants.py:9,35-43)."""
DEFAULT_SCHEDULING_UNIT = 100
INTER_TRAIN_PROCESS_COMM_PORT_1 = 12360
INTER_TRAIN_PROCESS_COMM_PORT_2 = 12376
