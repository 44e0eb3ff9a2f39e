"""Control-plane messaging between the ranks of one trial: a star of ZeroMQ sockets around a hub.

Same wire behaviour as the reference's ZMQBroadcastServer / ZMQBroadcastClient
(reference harness/determined/ipc.py:34-236): hub -> spokes over PUB/SUB, spokes -> hub over PUSH/PULL,
a hello handshake that defeats the PUB/SUB slow-joiner race, and a serial number on every message so
a rank that skips or repeats a collective is detected immediately (RuntimeError).

Only small pickled Python objects travel here (the ncclUniqueId, metric dictionaries, flags); tensor
payloads use the NCCL communicator of libdgs.so.
"""
import pickle
from typing import Any, List, Optional


class _Hello:
    final = False


class _LastHello(_Hello):
    final = True


class _Numbered:
    __slots__ = ("serial", "payload")

    def __init__(self, serial: int, payload: Any) -> None:
        self.serial, self.payload = serial, payload


def _expect_numbered(obj: Any, want: int, who: str) -> Any:
    if not isinstance(obj, _Numbered):
        raise RuntimeError(f"non-_SerialMessage: {type(obj).__name__}")
    if obj.serial != want:
        raise RuntimeError(f"Out-of-order {who} message detected: {obj.serial} != {want}")
    return obj.payload


class ZMQBroadcastServer:
    """The hub: broadcast() to every spoke, gather() one message from each."""

    def __init__(self, num_connections: int, pub_url: Optional[str] = None, pull_url: Optional[str] = None) -> None:
        import zmq

        self._n = num_connections
        self._ctx = zmq.Context()
        self._pub = self._ctx.socket(zmq.PUB)
        self._pull = self._ctx.socket(zmq.PULL)
        self._pub_port = self._bind(self._pub, pub_url)
        self._pull_port = self._bind(self._pull, pull_url)
        self._sent = 0
        self._received = 0

    @staticmethod
    def _bind(sock: Any, url: Optional[str]) -> Optional[int]:
        if url is None:
            return int(sock.bind_to_random_port("tcp://*"))
        sock.bind(url)
        return None

    def safe_start(self) -> None:
        """Keep publishing hellos until every spoke has answered one, then publish the last hello."""
        answered = 0
        while answered < self._n:
            self._pub.send(pickle.dumps(_Hello()))
            if not self._pull.poll(50):
                continue
            msg = pickle.loads(self._pull.recv())
            if not isinstance(msg, _Hello):
                raise RuntimeError(f"got non-_HelloMessage: {type(msg).__name__}")
            answered += 1
        self._pub.send(pickle.dumps(_LastHello()))

    def get_pub_port(self) -> int:
        if self._pub_port is None:
            raise ValueError("get_pub_port() is only safe when pub_url was None")
        return self._pub_port

    def get_pull_port(self) -> int:
        if self._pull_port is None:
            raise ValueError("get_pull_port() is only safe when pull_url was None")
        return self._pull_port

    def broadcast(self, obj: Any) -> None:
        self._pub.send(pickle.dumps(_Numbered(self._sent, obj)))
        self._sent += 1

    def gather(self) -> List[Any]:
        got = [_expect_numbered(pickle.loads(self._pull.recv()), self._received, "client") for _ in range(self._n)]
        self._received += 1
        return got

    def close(self) -> None:
        self._pub.close()
        self._pull.close()

    def __enter__(self) -> "ZMQBroadcastServer":
        return self

    def __exit__(self, *_: Any) -> None:
        self.close()


class ZMQBroadcastClient:
    """A spoke: send() to the hub, recv() the hub's next broadcast."""

    def __init__(self, srv_pub_url: str, srv_pull_url: str) -> None:
        import zmq

        self._ctx = zmq.Context()
        self._sub = self._ctx.socket(zmq.SUB)
        self._sub.subscribe(b"")
        self._sub.connect(srv_pub_url)
        self._push = self._ctx.socket(zmq.PUSH)
        self._push.connect(srv_pull_url)
        self._sent = 0
        self._received = 0

    def _next_hello(self) -> _Hello:
        msg = pickle.loads(self._sub.recv())
        if not isinstance(msg, _Hello):
            raise RuntimeError(f"got non-_HelloMessage: {type(msg).__name__}")
        return msg

    def safe_start(self) -> None:
        self._next_hello()  # the SUB socket is now known to be connected
        self._push.send(pickle.dumps(_Hello()))
        while not self._next_hello().final:
            pass

    def send(self, obj: Any) -> None:
        self._push.send(pickle.dumps(_Numbered(self._sent, obj)))
        self._sent += 1

    def recv(self) -> Any:
        payload = _expect_numbered(pickle.loads(self._sub.recv()), self._received, "server")
        self._received += 1
        return payload

    def close(self) -> None:
        self._sub.close()
        self._push.close()

    def __enter__(self) -> "ZMQBroadcastClient":
        return self

    def __exit__(self, *_: Any) -> None:
        self.close()
