"""Shared queue: the put/get/size contract of the reference's Ray actor, without Ray.

Reference: ``psana_ray/shared_queue.py`` -- ``@ray.remote class Queue`` around ``deque(maxlen)``
(:4-31) registered by ``create_queue`` as a *named, namespaced, detached* actor (:33-38).  Ray is
not installable here, so the same contract is served by a small single-threaded server process on
a unix-domain socket (``queue_server.py``): one operation at a time, FIFO across all producers and
consumers, found by ``(queue_name, namespace)`` from any process on the box, outliving its creator.

Kept verbatim from the reference:
  * ``put`` never blocks and never evicts: ``True`` if there was room, else ``False`` (:11-14);
  * ``get`` never blocks: the oldest item, or ``None`` when empty (:21);
  * ``size`` (:28);  ``None`` is a storable item (the producer's end-of-stream sentinel, producer.py:125);
  * errors inside an operation are swallowed and give ``False`` / ``None`` / ``0`` (:15-17,:22-24,:29-31);
  * a dead queue is NOT swallowed: the handle raises ``QueueDeadError`` (Ray: ``RayActorError``).

What changes is the payload: a calibrated batch stays on the GPU and travels as a 64-byte CUDA-IPC
handle (``transport.GpuBatchRef``) instead of a pickled ndarray through the object store.  A batch
item carries ``weight`` = its number of events so that ``maxsize`` keeps counting events.
"""
from __future__ import annotations

import os
import pickle
import socket
import struct
import subprocess
import sys
import time
from collections import deque

DEFAULT_ROOT_ENV = "PSANA_RAY_DIR"


class QueueDeadError(RuntimeError):
    """The queue server is gone (the analogue of ray.exceptions.RayActorError)."""


class QueueNotFoundError(ValueError):
    """No queue of that name in that namespace (Ray raises ValueError from get_actor)."""


# ---------------------------------------------------------------------------------------------------
# the in-process queue: exactly the reference actor's body (ref: shared_queue.py:4-31), plus weights
# ---------------------------------------------------------------------------------------------------
class Queue:
    def __init__(self, maxsize=100):
        self.items = deque(maxlen=None)       # length is enforced in put(); the reference's maxlen never evicts either
        self.maxlen = maxsize
        self.weight = 0

    def put(self, item, weight=1):
        try:
            # ref :11 `len < maxlen`; with weighted (batch) items an empty queue always accepts one item
            if self.weight < self.maxlen and (self.weight + weight <= self.maxlen or not self.items):
                self.items.append((item, weight))
                self.weight += weight
                return True
            return False
        except Exception as e:
            print(f"Error in put: {e}")
            return False

    def get(self):
        try:
            if not self.items:
                return None
            item, w = self.items.popleft()
            self.weight -= w
            return item
        except Exception as e:
            print(f"Error in get: {e}")
            return None

    def get_ex(self):
        """(found, item): lets a caller tell an empty queue from a stored ``None`` sentinel."""
        if not self.items:
            return False, None
        item, w = self.items.popleft()
        self.weight -= w
        return True, item

    def size(self):
        try:
            return self.weight
        except Exception as e:
            print(f"Error in size: {e}")
            return 0


# ---------------------------------------------------------------------------------------------------
# wire format: 4-byte big-endian length + pickle
# ---------------------------------------------------------------------------------------------------
def send_msg(sock, obj):
    data = pickle.dumps(obj, protocol=5)
    sock.sendall(struct.pack(">I", len(data)) + data)


def recv_exact(sock, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(n - len(buf))
        if not chunk:
            raise ConnectionError("peer closed")
        buf += chunk
    return bytes(buf)


def recv_msg(sock):
    (n,) = struct.unpack(">I", recv_exact(sock, 4))
    return pickle.loads(recv_exact(sock, n))


# ---------------------------------------------------------------------------------------------------
# discovery: (address, namespace, name) -> socket path
# ---------------------------------------------------------------------------------------------------
def resolve_root(address="auto"):
    """'auto' -> $PSANA_RAY_DIR or /tmp/psana_ray_<uid>; anything else is taken as a directory."""
    if address in (None, "auto", ""):
        return os.environ.get(DEFAULT_ROOT_ENV) or os.path.join("/tmp", f"psana_ray_{os.getuid()}")
    if address.startswith("unix://"):
        return address[len("unix://"):]
    return address


def socket_path(queue_name, namespace, address="auto"):
    return os.path.join(resolve_root(address), namespace, queue_name + ".sock")


class QueueHandle:
    """Client side of one named queue.  One connection per handle; not thread-safe (like a Ray handle used
    from one driver thread)."""

    def __init__(self, path):
        self.path = path
        self._sock = None
        self._connect()

    def _connect(self):
        s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        try:
            s.connect(self.path)
        except OSError as e:
            s.close()
            raise QueueNotFoundError(f"no queue at {self.path}: {e}") from e
        self._sock = s

    def _call(self, op, *args):
        if self._sock is None:
            raise QueueDeadError("handle is closed")
        try:
            send_msg(self._sock, (op,) + args)
            ok, val = recv_msg(self._sock)
        except (OSError, ConnectionError, EOFError, struct.error) as e:
            raise QueueDeadError(f"Queue server at {self.path} is dead: {e}") from e
        if not ok:
            raise RuntimeError(val)
        return val

    # -- the reference contract ------------------------------------------------------------------------
    def put(self, item, weight=1, lease=None):
        return bool(self._call("put", pickle.dumps(item, protocol=5), int(weight), lease))

    def get(self):
        found, blob = self._call("get")
        return pickle.loads(blob) if found and blob is not None else None

    def size(self):
        return int(self._call("size"))

    # -- additive ---------------------------------------------------------------------------------------
    def get_ex(self):
        """(found, item) -- distinguishes EMPTY from a stored None sentinel."""
        found, blob = self._call("get")
        return bool(found), (pickle.loads(blob) if found and blob is not None else None)

    def release(self, lease):
        """Consumer is done with a leased GPU slab; the producer may reuse it."""
        return self._call("release", lease)

    def poll_released(self, owner):
        """Producer side: leases of ``owner`` that have been released since the last poll."""
        return self._call("poll_released", owner)

    def register_slabs(self, table):
        """Producer side: publish the slab ring's handle table once (transport.SlabRing.table())."""
        return self._call("register_slabs", table)

    def list_slabs(self):
        """Consumer side: every registered producer's slab table."""
        return self._call("list_slabs")

    def barrier(self, tag, size, timeout=600.0, poll=0.01):
        """Host-side rendezvous standing in for MPI.COMM_WORLD.Barrier() (ref: producer.py:53,120)."""
        target = self._call("barrier_arrive", tag, int(size))
        t0 = time.time()
        while not self._call("barrier_done", tag, target):
            if time.time() - t0 > timeout:
                raise TimeoutError(f"barrier {tag!r} timed out")
            time.sleep(poll)

    def stats(self):
        return self._call("stats")

    def shutdown(self):
        """Stop the detached server (the analogue of ray.kill(actor))."""
        try:
            self._call("shutdown")
        except QueueDeadError:
            pass
        self.close()

    def close(self):
        if self._sock is not None:
            try:
                self._sock.close()
            finally:
                self._sock = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def get_queue(queue_name="shared_queue", ray_namespace="default", address="auto"):
    """``ray.get_actor(name, namespace)``: raises ``QueueNotFoundError`` (a ValueError) when absent."""
    path = socket_path(queue_name, ray_namespace, address)
    if not os.path.exists(path):
        raise QueueNotFoundError(f"Failed to look up queue {queue_name!r} in namespace {ray_namespace!r} ({path})")
    return QueueHandle(path)


def create_queue(queue_name="shared_queue", ray_namespace="default", maxsize=100, address="auto", wait_s=10.0):
    """Start a detached queue server and return a handle, or ``None`` on failure (ref: shared_queue.py:33-38)."""
    try:
        path = socket_path(queue_name, ray_namespace, address)
        # Messages are pickles, so whoever can connect can run code as the queue's owner (as with a Ray actor): the
        # directories this call creates are private to the user, and the server makes the socket 0600.
        d = os.path.dirname(path)
        missing = []
        while d and not os.path.isdir(d):
            missing.append(d)
            d = os.path.dirname(d)
        os.makedirs(os.path.dirname(path), mode=0o700, exist_ok=True)
        for m in missing:
            os.chmod(m, 0o700)
        if os.path.exists(path):
            try:
                QueueHandle(path).close()
                raise ValueError(f"queue {queue_name!r} already exists in namespace {ray_namespace!r}")   # Ray: name taken
            except QueueNotFoundError:
                os.unlink(path)                  # stale socket of a dead server
        env = dict(os.environ)
        pkg_parent = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        env["PYTHONPATH"] = pkg_parent + os.pathsep + env.get("PYTHONPATH", "")
        subprocess.Popen([sys.executable, "-m", "psana_ray.queue_server", "--path", path, "--maxsize", str(int(maxsize))],
                         env=env, start_new_session=True, stdin=subprocess.DEVNULL,
                         stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, close_fds=True)
        t0 = time.time()
        while time.time() - t0 < wait_s:
            if os.path.exists(path):
                try:
                    return QueueHandle(path)
                except QueueNotFoundError:
                    pass
            time.sleep(0.02)
        raise TimeoutError(f"queue server did not come up at {path}")
    except Exception as e:
        print(f"Error creating queue '{queue_name}' in namespace '{ray_namespace}': {e}")
        return None
