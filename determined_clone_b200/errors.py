"""Exception types of the reference harness used on the grad-sync path
(reference harness/determined/errors.py; raised at harness/determined/pytorch/_pytorch_context.py:289,347,351,715,869)."""


class InternalException(Exception):
    """The harness itself is in an unexpected state."""


class InvalidExperimentException(Exception):
    """The user's experiment or trial code misuses the API."""


class InvalidCheckpointException(Exception):
    """A checkpoint directory does not hold what its format promises (reference errors.py:45-51)."""

    def __init__(self) -> None:
        super().__init__("Invalid checkpoint.")
