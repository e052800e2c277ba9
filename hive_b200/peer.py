"""Device buffers shared between the ranks of one node (one process per GPU) through CUDA IPC -- the plumbing under the
peer-to-peer mesh assembly of ``hive_b200.sharded``: rank 0 owns the whole-mesh arrays, every slab's emit kernel stores its
rows into them over NVLink (``hv_mc_emit_direct``).  Nothing in the reference corresponds to this.

The owner allocates with ``hv_peer_alloc`` (a plain ``cudaMalloc``: an IPC handle covers a whole allocation, so a block of
torch's caching allocator will not do), broadcasts the 64-byte handle over the process group the ranks already have, and the
others map it with ``hv_peer_open``.  Whether that worked is agreed on by all ranks (one all-reduce), so either every rank uses
the shared buffer or none does.
"""
from __future__ import annotations

import ctypes

from . import _lib

HANDLE_BYTES = 64
_ALIGN = 256


def mesh_layout(n_verts: int, n_tris: int):
    """Byte offsets of the whole mesh's arrays inside the shared buffer -> ({name: offset}, total bytes).  Every rank
    computes this from the all-gathered totals, so nothing about the layout is communicated."""
    offs, o = {}, 0
    for name, nbytes in (("verts", 12 * n_verts), ("normals", 12 * n_verts), ("colors", 3 * n_verts), ("faces", 12 * n_tris)):
        offs[name] = o
        o = (o + nbytes + _ALIGN - 1) & ~(_ALIGN - 1)
    return offs, max(o, _ALIGN)


class SharedBuffer:
    """``nbytes`` of device memory on ``owner``'s GPU, addressable from every rank of ``dist``'s default group.

    Collective: every rank constructs it at the same point.  ``self.ok`` is the same on every rank; ``self.base`` is this
    process's address of the buffer (the owner's own pointer, a mapping elsewhere).
    """

    def __init__(self, dist, rank: int, device, nbytes: int, owner: int = 0):
        torch = _lib.require_cuda()
        lib = _lib.load()
        self._lib, self._dist, self.rank, self.owner, self.device, self.nbytes = lib, dist, rank, owner, device, int(nbytes)
        self.base, self.ok = None, False
        good = 1
        msg = torch.zeros(HANDLE_BYTES + 8, dtype=torch.uint8, device=device)
        with torch.cuda.device(device):
            if rank == owner:
                ptr, handle = ctypes.c_void_p(0), (ctypes.c_uint8 * HANDLE_BYTES)()
                if lib.hv_peer_alloc(self.nbytes, ctypes.byref(ptr), handle) == 0:
                    self.base = int(ptr.value)
                    msg[:HANDLE_BYTES] = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(device)
                    msg[HANDLE_BYTES] = 1
                else:
                    good = 0
            dist.broadcast(msg, src=owner)
            if rank != owner:
                host = msg.cpu().numpy()
                if host[HANDLE_BYTES] == 1:
                    ptr = ctypes.c_void_p(0)
                    handle = (ctypes.c_uint8 * HANDLE_BYTES).from_buffer_copy(host[:HANDLE_BYTES].tobytes())
                    if lib.hv_peer_open(handle, ctypes.byref(ptr)) == 0:
                        self.base = int(ptr.value)
                    else:
                        good = 0
                else:
                    good = 0
            flag = torch.tensor([good], dtype=torch.int32, device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            self.ok = bool(int(flag.item()) == 1)
        if not self.ok:
            self.error = _lib.last_error()
            self.close()

    def close(self):
        """Collective: the mappings go first, then the owner frees."""
        if self.base is not None and self.rank != self.owner:
            self._lib.hv_peer_close(ctypes.c_void_p(self.base))
            self.base = None
        self._dist.barrier()
        if self.base is not None:
            self._lib.hv_peer_free(ctypes.c_void_p(self.base))
            self.base = None

    def copy_out(self, offset: int, host_tensor, stream):
        """Owner: asynchronous copy of ``host_tensor.nbytes`` bytes at ``offset`` into a pinned host tensor."""
        n = host_tensor.numel() * host_tensor.element_size()
        assert self.base is not None and offset + n <= self.nbytes
        _lib.check(self._lib.hv_copy_async(ctypes.c_void_p(host_tensor.data_ptr()), ctypes.c_void_p(self.base + offset), n,
                                           ctypes.c_void_p(stream)))


class _RawDevice:
    """A device address as something ``torch.as_tensor`` accepts (``__cuda_array_interface__``): no ownership."""

    def __init__(self, address: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(address), False), "version": 3}


def view(address: int, shape, dtype, device):
    """A torch tensor over device memory torch does not own (a buffer of ``hv_peer_alloc`` / ``hv_peer_open``)."""
    torch = _lib.require_cuda()
    n = int(torch.empty((), dtype=dtype).element_size())
    for s in shape:
        n *= int(s)
    return torch.as_tensor(_RawDevice(address, n), device=device).view(dtype).view(*shape)


class Mailboxes:
    """One buffer per rank, every one mapped into every rank: what the slabs of a sharded volume write to each other
    directly (seam planes, arrival flags) or read from each other (per-column counts) instead of going through a collective.

    ``self.base[w]`` is this process's address of rank w's buffer (``self.base[rank]`` its own).  Collective: every rank
    constructs it at the same point; ``self.ok`` is the same everywhere.
    """

    def __init__(self, dist, rank: int, world: int, device, nbytes: int):
        torch = _lib.require_cuda()
        lib = _lib.load()
        self._lib, self._dist, self.rank, self.world, self.device, self.nbytes = lib, dist, rank, world, device, int(nbytes)
        self.base, self.ok, self._own = [None] * world, False, None
        good = 1
        with torch.cuda.device(device):
            mine = torch.zeros(HANDLE_BYTES + 8, dtype=torch.uint8, device=device)
            ptr, handle = ctypes.c_void_p(0), (ctypes.c_uint8 * HANDLE_BYTES)()
            if lib.hv_peer_alloc(self.nbytes, ctypes.byref(ptr), handle) == 0:
                self._own = self.base[rank] = int(ptr.value)
                view(self._own, (self.nbytes,), torch.uint8, device).zero_()  # flags start at epoch 0
                mine[:HANDLE_BYTES] = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(device)
                mine[HANDLE_BYTES] = 1
            else:
                good = 0
            torch.cuda.current_stream().synchronize()
            every = torch.empty((world, HANDLE_BYTES + 8), dtype=torch.uint8, device=device)
            dist.all_gather_into_tensor(every.view(-1), mine)
            host = every.cpu().numpy()
            for w in range(world):
                if w == rank:
                    continue
                if host[w, HANDLE_BYTES] != 1:
                    good = 0
                    continue
                p = ctypes.c_void_p(0)
                h = (ctypes.c_uint8 * HANDLE_BYTES).from_buffer_copy(host[w, :HANDLE_BYTES].tobytes())
                if lib.hv_peer_open(h, ctypes.byref(p)) == 0:
                    self.base[w] = int(p.value)
                else:
                    good = 0
            flag = torch.tensor([good], dtype=torch.int32, device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            self.ok = bool(int(flag.item()) == 1)
        if not self.ok:
            self.error = _lib.last_error()
            self.close()

    def close(self):
        """Collective: the mappings go first, then every rank frees its own buffer."""
        for w in range(self.world):
            if w != self.rank and self.base[w] is not None:
                self._lib.hv_peer_close(ctypes.c_void_p(self.base[w]))
                self.base[w] = None
        self._dist.barrier()
        if self._own is not None:
            self._lib.hv_peer_free(ctypes.c_void_p(self._own))
            self._own = self.base[self.rank] = None

    def signal(self, addresses, epoch: int, stream: int):
        """Store ``epoch`` into the flags at ``addresses`` (this process's addresses of words in any rank's buffer), after
        everything queued on ``stream`` so far."""
        arr = (ctypes.c_void_p * len(addresses))(*[ctypes.c_void_p(a) for a in addresses])
        _lib.check(self._lib.hv_peer_signal(arr, len(addresses), ctypes.c_uint(epoch & 0xFFFFFFFF), ctypes.c_void_p(stream)))

    def wait(self, offset: int, n: int, epoch: int, err, stream: int):
        """Hold ``stream`` until the ``n`` consecutive flags at ``offset`` of this rank's own buffer have reached ``epoch``."""
        _lib.check(self._lib.hv_peer_wait(ctypes.c_void_p(self._own + offset), n, ctypes.c_uint(epoch & 0xFFFFFFFF),
                                          ctypes.c_void_p(err.data_ptr()), ctypes.c_void_p(stream)))
