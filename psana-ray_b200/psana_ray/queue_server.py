"""Detached queue server: one named queue on one unix-domain socket.

Stands in for the Ray actor process of the reference (ref: psana_ray/shared_queue.py:4-35 --
``Queue.options(name=..., namespace=..., lifetime="detached").remote(maxsize)``): a single thread
serves one request at a time, so every put/get of every rank and consumer is serialised here and
the queue is FIFO in arrival order, exactly like a default (max_concurrency=1) Ray actor.

Items are opaque pickled blobs (the server never unpickles them and never imports torch or CUDA);
GPU batches are 64-byte CUDA-IPC descriptors, so the server moves bytes, not frames.

Beyond put/get/size the server keeps the bookkeeping the GPU hand-off needs:
  * leases -- a batch item may name the producer slab it lives in; the lease is attached to the
    consumer connection that got it and is released explicitly, or automatically when that
    connection drops (a crashed consumer must not wedge the producer's ring);
  * a counting barrier, standing in for MPI.COMM_WORLD.Barrier() (ref: producer.py:53,120);
  * the producers' slab tables (CUDA-IPC memory and event handles), so that a consumer can map every
    slab when it connects instead of on first touch inside the stream.

    python -m psana_ray.queue_server --path /tmp/psana_ray_0/default/my.sock --maxsize 400
"""
from __future__ import annotations

import argparse
import os
import selectors
import socket
import struct
import sys

from .shared_queue import Queue


class Conn:
    __slots__ = ("sock", "buf", "leases")

    def __init__(self, sock):
        self.sock = sock
        self.buf = bytearray()
        self.leases = set()


class QueueServer:
    def __init__(self, path, maxsize):
        self.path = path
        self.queue = Queue(maxsize)
        self.released = {}          # owner -> [lease, ...] waiting to be polled by the producer
        self.barriers = {}          # tag -> arrivals so far
        self.slab_tables = {}       # owner -> table of a producer's slab ring (memory + event handles), registered once
        self.nput = self.nget = self.nrefused = 0
        self.running = True
        self.sel = selectors.DefaultSelector()
        if os.path.exists(path):
            os.unlink(path)
        self.lsock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self.lsock.bind(path + ".tmp")
        os.chmod(path + ".tmp", 0o600)             # requests are pickles: only the owner may connect
        self.lsock.listen(256)
        os.rename(path + ".tmp", path)          # the name appears only once the server is listening
        self.lsock.setblocking(False)
        self.sel.register(self.lsock, selectors.EVENT_READ, None)

    # -- operations (each runs alone; errors are swallowed as in the reference actor) -------------------
    def op_put(self, conn, blob, weight, lease):
        try:
            ok = self.queue.put((blob, lease), weight)
            if ok:
                self.nput += 1
            else:
                self.nrefused += 1
            return ok
        except Exception as e:
            print(f"Error in put: {e}")
            return False

    def op_get(self, conn):
        try:
            found, item = self.queue.get_ex()
            if not found:
                return (False, None)
            blob, lease = item
            if lease is not None:
                conn.leases.add(lease)
            self.nget += 1
            return (True, blob)
        except Exception as e:
            print(f"Error in get: {e}")
            return (False, None)

    def op_size(self, conn):
        return self.queue.size()

    def _release(self, lease):
        self.released.setdefault(lease[0], []).append(lease)

    def op_release(self, conn, lease):
        lease = tuple(lease)
        conn.leases.discard(lease)
        self._release(lease)
        return True

    def op_poll_released(self, conn, owner):
        return self.released.pop(owner, [])

    def op_register_slabs(self, conn, table):
        self.slab_tables[table["owner"]] = table
        return True

    def op_list_slabs(self, conn):
        return list(self.slab_tables.values())

    def op_barrier_arrive(self, conn, tag, size):
        """Returns the arrival count this caller must wait for (a tag can be reused run after run)."""
        n = self.barriers.get(tag, 0) + 1
        self.barriers[tag] = n
        return -(-n // size) * size

    def op_barrier_done(self, conn, tag, target):
        return self.barriers.get(tag, 0) >= target

    def op_stats(self, conn):
        return {"size": self.queue.size(), "items": len(self.queue.items), "maxsize": self.queue.maxlen,
                "put": self.nput, "get": self.nget, "refused": self.nrefused, "pid": os.getpid(),
                "leases_out": sum(len(c.leases) for c in self._conns())}

    def op_shutdown(self, conn):
        self.running = False
        return True

    def _conns(self):
        return [k.data for k in self.sel.get_map().values() if k.data is not None]

    # -- event loop -------------------------------------------------------------------------------------
    def _drop(self, conn):
        for lease in list(conn.leases):      # consumer went away holding slabs: give them back
            self._release(lease)
        conn.leases.clear()
        try:
            self.sel.unregister(conn.sock)
        except Exception:
            pass
        conn.sock.close()

    def _serve(self, conn):
        import pickle
        try:
            chunk = conn.sock.recv(1 << 20)
        except BlockingIOError:
            return
        except OSError:
            chunk = b""
        if not chunk:
            self._drop(conn)
            return
        conn.buf += chunk
        while len(conn.buf) >= 4:
            (n,) = struct.unpack(">I", conn.buf[:4])
            if len(conn.buf) < 4 + n:
                break
            msg = pickle.loads(bytes(conn.buf[4:4 + n]))
            del conn.buf[:4 + n]
            try:
                fn = getattr(self, "op_" + msg[0])
                reply = (True, fn(conn, *msg[1:]))
            except Exception as e:
                reply = (False, f"{type(e).__name__}: {e}")
            data = pickle.dumps(reply, protocol=5)
            try:
                conn.sock.setblocking(True)
                conn.sock.sendall(struct.pack(">I", len(data)) + data)
                conn.sock.setblocking(False)
            except OSError:
                self._drop(conn)
                return

    def run(self):
        try:
            while self.running:
                for key, _ in self.sel.select(timeout=1.0):
                    if key.data is None:
                        try:
                            s, _ = self.lsock.accept()
                        except OSError:
                            continue
                        s.setblocking(False)
                        self.sel.register(s, selectors.EVENT_READ, Conn(s))
                    else:
                        self._serve(key.data)
        finally:
            for c in self._conns():
                c.sock.close()
            self.lsock.close()
            try:
                os.unlink(self.path)
            except OSError:
                pass


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--path", required=True)
    ap.add_argument("--maxsize", type=int, default=100)
    a = ap.parse_args(argv)
    QueueServer(a.path, a.maxsize).run()


if __name__ == "__main__":
    sys.exit(main())
