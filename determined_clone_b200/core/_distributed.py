"""DistributedContext: rank bookkeeping plus gather / allgather / broadcast of Python objects.

Drop-in for the reference's determined.core.DistributedContext
(reference harness/determined/core/_distributed.py:12-419): same constructor, class methods,
method names, return shapes (results ordered by rank, workers get None from gather) and errors.
Objects still travel over ZeroMQ through the chief (ipc.py).  New here: once a libdgs
communicator is attached (attach_comm), *tensor* payloads can use NCCL over NVLink instead --
allgather_tensor / broadcast_tensor -- which is what the metric reducer and the parameter
broadcast use.
"""
import logging
import os
import shutil
import socket
import tempfile
from typing import Any, Callable, List, Optional

from determined_clone_b200 import constants, ipc

logger = logging.getLogger("determined.core")


class _Star:
    """One hub-and-spokes group (the whole job, or the ranks of one machine)."""

    def __init__(self, index: int, size: int) -> None:
        self.index, self.size = index, size
        self.hub: Optional[ipc.ZMQBroadcastServer] = None
        self.spoke: Optional[ipc.ZMQBroadcastClient] = None

    @property
    def is_hub(self) -> bool:
        return self.index == 0

    def gather(self, stuff: Any, with_barrier: bool) -> Optional[List[Any]]:
        if self.size < 2:
            return [stuff]
        if self.hub is not None:
            ranked = self.hub.gather()
            out = [stuff] + [value for _, value in sorted(ranked, key=lambda rv: rv[0])]
            if with_barrier:
                # workers must not start the next gather before everyone has finished this one
                self.hub.broadcast(None)
            return out
        assert self.spoke is not None
        self.spoke.send((self.index, stuff))
        if with_barrier:
            self.spoke.recv()
        return None

    def allgather(self, stuff: Any) -> List[Any]:
        if self.size < 2:
            return [stuff]
        if self.hub is not None:
            everything = self.gather(stuff, with_barrier=False)
            self.hub.broadcast(everything)
            return everything  # type: ignore
        assert self.spoke is not None
        self.spoke.send((self.index, stuff))
        return self.spoke.recv()  # type: ignore

    def broadcast(self, stuff: Any) -> Any:
        if self.size < 2:
            return stuff
        if self.hub is not None:
            self.hub.broadcast(stuff)
            return stuff
        assert self.spoke is not None
        return self.spoke.recv()

    def close(self) -> None:
        for end in (self.hub, self.spoke):
            if end is not None:
                end.close()
        self.hub = self.spoke = None


class DistributedContext:
    """Rank information and small-object collectives for the workers of one trial.

    Required keyword arguments: rank, size, local_rank, local_size, cross_rank, cross_size
    (all six or none).  chief_ip is required when the job spans machines.
    """

    def __init__(
        self,
        *,
        rank: int,
        size: int,
        local_rank: int,
        local_size: int,
        cross_rank: int,
        cross_size: int,
        chief_ip: Optional[str] = None,
        pub_port: int = constants.INTER_TRAIN_PROCESS_COMM_PORT_1,
        pull_port: int = constants.INTER_TRAIN_PROCESS_COMM_PORT_2,
        port_offset: int = 0,
        force_tcp: bool = False,
    ) -> None:
        given = [x is not None for x in (rank, size, local_rank, local_size, cross_rank, cross_size)]
        if any(given) and not all(given):
            raise ValueError(
                "rank, size, local_rank, local_size, cross_rank, and cross_size must all be "
                "provided if any are provided"
            )
        self.rank = rank if rank is not None else 0
        self.size = size if size is not None else 1
        self.local_rank = local_rank if local_rank is not None else 0
        self.local_size = local_size if local_size is not None else 1
        self.cross_rank = cross_rank if cross_rank is not None else 0
        self.cross_size = cross_size if cross_size is not None else 1

        self._pub_port = pub_port + port_offset
        self._pull_port = pull_port + port_offset
        self._is_chief = self.rank == 0
        self._is_local_chief = self.local_rank == 0
        if self.local_size != self.size:
            if chief_ip is None:
                raise AssertionError(
                    f"This is a distributed job (local_size {self.local_size} != size {self.size}) "
                    "but chief_ip was none provided.  When cross_size > 1, the chief_ip parameter "
                    "is required."
                )
            self._chief_ip = chief_ip
        else:
            self._chief_ip = "127.0.0.1"  # single machine: always reach the chief over loopback

        self._closed = False
        self.tempdir: Optional[str] = None
        self._global = _Star(self.rank, self.size)
        self._local = _Star(self.local_rank, self.local_size)
        self._comm: Any = None  # libdgs communicator wrapper, see attach_comm()
        self._init_ipc(force_tcp)

    # ---- set-up / tear-down -----------------------------------------------------------------
    def _init_ipc(self, force_tcp: bool) -> None:
        if self.size < 2:
            return
        if self._is_chief:
            self._global.hub = ipc.ZMQBroadcastServer(
                num_connections=self.size - 1,
                pub_url=f"tcp://*:{self._pub_port}",
                pull_url=f"tcp://*:{self._pull_port}",
            )
            self._global.hub.safe_start()
        else:
            self._global.spoke = ipc.ZMQBroadcastClient(
                srv_pub_url=f"tcp://{self._chief_ip}:{self._pub_port}",
                srv_pull_url=f"tcp://{self._chief_ip}:{self._pull_port}",
            )
            self._global.spoke.safe_start()

        # Every rank takes part in exactly one global allgather here, through which each machine's
        # local chief announces where its local star listens.
        if self.local_size < 2:
            self.allgather(None)
            return
        if self._is_local_chief:
            pub_url = pull_url = None
            if hasattr(socket, "AF_UNIX") and not force_tcp:
                self.tempdir = tempfile.mkdtemp(prefix="ipc")
                pub_url = f"ipc://{self.tempdir}/pub.sock"
                pull_url = f"ipc://{self.tempdir}/pull.sock"
            self._local.hub = ipc.ZMQBroadcastServer(self.local_size - 1, pub_url=pub_url, pull_url=pull_url)
            if pub_url is None:
                pub_url = f"tcp://localhost:{self._local.hub.get_pub_port()}"
            if pull_url is None:
                pull_url = f"tcp://localhost:{self._local.hub.get_pull_port()}"
            self.allgather((self.cross_rank, pub_url, pull_url))
            self._local.hub.safe_start()
        else:
            announced = [x for x in self.allgather(None) if x is not None and x[0] == self.cross_rank]
            assert len(announced) == 1, (
                f"did not find exactly 1 local_chief for machine {self.cross_rank} in {announced}"
            )
            _, pub_url, pull_url = announced[0]
            assert isinstance(pub_url, str), f"invalid pub_url: {pub_url}"
            assert isinstance(pull_url, str), f"invalid pub_url: {pull_url}"
            self._local.spoke = ipc.ZMQBroadcastClient(pub_url, pull_url)
            self._local.spoke.safe_start()

    def close(self) -> None:
        if self._closed or self.size < 2:
            return
        self._closed = True
        if self._comm is not None:
            self._comm.close()
            self._comm = None
        self._global.close()
        self._local.close()
        if self.tempdir is not None:
            shutil.rmtree(self.tempdir, ignore_errors=True)
            self.tempdir = None

    # ---- constructors from launcher environments ------------------------------------------------
    @classmethod
    def from_horovod(cls, hvd: Any, chief_ip: Optional[str] = None) -> "DistributedContext":
        return cls(
            rank=hvd.rank(),
            size=hvd.size(),
            local_rank=hvd.local_rank(),
            local_size=hvd.local_size(),
            cross_rank=hvd.cross_rank(),
            cross_size=hvd.cross_size(),
            chief_ip=chief_ip or os.environ.get("DET_CHIEF_IP"),
        )

    @classmethod
    def _from_env(cls, names: List[str], chief_ip: Optional[str]) -> "DistributedContext":
        rank, size, local_rank, local_size, cross_rank, cross_size = (int(os.environ[n]) for n in names)
        return cls(
            rank=rank, size=size, local_rank=local_rank, local_size=local_size, cross_rank=cross_rank,
            cross_size=cross_size, chief_ip=chief_ip or os.environ.get("DET_CHIEF_IP"),
        )

    @classmethod
    def from_deepspeed(cls, chief_ip: Optional[str] = None) -> "DistributedContext":
        return cls._from_env(["RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_SIZE", "CROSS_RANK", "CROSS_SIZE"], chief_ip)

    @classmethod
    def from_torch_distributed(cls, chief_ip: Optional[str] = None) -> "DistributedContext":
        return cls._from_env(
            ["RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE", "GROUP_RANK", "GROUP_WORLD_SIZE"], chief_ip
        )

    # ---- rank queries -------------------------------------------------------------------------
    def get_rank(self) -> int:
        return self.rank

    def get_local_rank(self) -> int:
        return self.local_rank

    def get_size(self) -> int:
        return self.size

    def get_num_agents(self) -> int:
        return self.cross_size

    # ---- object collectives (ZeroMQ through the chief) ------------------------------------------
    def gather(self, stuff: Any) -> Optional[List]:
        """Chief gets [own, *others by rank]; workers get None.  Ends with a barrier."""
        return self._global.gather(stuff, with_barrier=True)

    def gather_local(self, stuff: Any) -> Optional[List]:
        return self._local.gather(stuff, with_barrier=True)

    def allgather(self, stuff: Any) -> List:
        """Everyone gets the list of everyone's stuff, ordered by rank."""
        return self._global.allgather(stuff)

    def allgather_local(self, stuff: Any) -> List:
        return self._local.allgather(stuff)

    def broadcast(self, stuff: Any) -> Any:
        """Everyone gets the chief's stuff."""
        return self._global.broadcast(stuff)

    def broadcast_local(self, stuff: Any = None) -> Any:
        return self._local.broadcast(stuff)

    # ---- tensor collectives (NCCL through libdgs) -----------------------------------------------
    def attach_comm(self, make_comm: Callable[["DistributedContext"], Any]) -> Any:
        """Create (once) the NCCL communicator for this job; the unique id travels over broadcast()."""
        if self._comm is None:
            self._comm = make_comm(self)
        return self._comm

    @property
    def comm(self) -> Any:
        return self._comm


class DummyDistributedContext(DistributedContext):
    def __init__(self) -> None:
        super().__init__(rank=0, size=1, local_rank=0, local_size=1, cross_rank=0, cross_size=1)
