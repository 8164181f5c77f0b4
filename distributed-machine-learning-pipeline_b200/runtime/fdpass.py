"""Passing a POSIX file descriptor from rank 0 to every other rank of a torch.distributed group.

Used for the NVLS multicast object (``cuMemExportToShareableHandle`` yields an fd that peers must import,
``runtime/nvls.py``). torch.distributed only carries the rendezvous (a Unix socket path); the descriptor itself
travels as ``SCM_RIGHTS`` ancillary data. Works with any backend (tested on CPU with gloo).
"""
import os
import socket
import tempfile
import time
from typing import Optional


def broadcast_fd(fd: Optional[int], group=None, tag: str = "dmlp", error: Optional[str] = None) -> int:
    """Rank 0 passes ``fd`` (or ``error``: a message that makes every rank raise RuntimeError). Returns the caller's
    descriptor: rank 0 its own, every other rank a new descriptor referring to the same open file description
    (the caller closes it). Collective: every rank of ``group`` must call it."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    srv, path = None, None
    if rank == 0 and error is None:
        path = os.path.join(tempfile.gettempdir(), f"{tag}_{os.getpid()}_{int(time.time() * 1e6)}.sock")
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(path)
        srv.listen(world)
    box = [error or "ok", path]
    dist.broadcast_object_list(box, src=0, group=group)
    if box[0] != "ok":
        raise RuntimeError(box[0])
    if rank == 0:
        try:
            for _ in range(world - 1):
                conn, _ = srv.accept()
                socket.send_fds(conn, [tag.encode()], [fd])
                conn.recv(1)  # the peer holds its own descriptor now
                conn.close()
        finally:
            srv.close()
            os.unlink(path)
        return fd
    c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    for _ in range(500):
        try:
            c.connect(box[1])
            break
        except (FileNotFoundError, ConnectionRefusedError):
            time.sleep(0.01)
    else:
        raise RuntimeError(f"could not reach rank 0's descriptor socket {box[1]}")
    _, fds, _, _ = socket.recv_fds(c, 64, 1)
    c.send(b"k")
    c.close()
    if not fds:
        raise RuntimeError("rank 0 sent no descriptor")
    return fds[0]
