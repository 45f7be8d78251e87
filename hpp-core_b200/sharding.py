"""Multi-GPU plumbing for the batched validation path.

One process per GPU (torch.distributed; NCCL on the GPU box, gloo in the CPU tests).  Configurations
and edges are independent units, so the path shards by contiguous index ranges with the model tables
replicated; the only exchange is the gather of the per-unit results (validity masks, 1 B per unit, and
optionally pair indices / parameters).  k-NN shards the NODE matrix by node index and replicates the
queries: every rank returns its local top-k as (distance, GLOBAL node index) and the candidates are
merged with the reference's tie rule -- strict `<` in nearestNeighbor::Basic keeps the EARLIER node on
equal distance (reference src/nearest-neighbor/basic.cc:66,106), i.e. the key is (distance, index).

Nothing here computes a distance or a validity flag: those come from the C ABI on each rank.
"""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous index range [begin, end) of `rank`: floor(rank*n/world) .. floor((rank+1)*n/world)."""
    return (rank * n) // world, ((rank + 1) * n) // world


def shard_sizes(n, world):
    return [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]


def node_owner_ranges(n_nodes, world):
    """Node-matrix shards for k-NN: rank r owns global node indices [begin_r, end_r)."""
    return [shard_range(n_nodes, r, world) for r in range(world)]


def all_gather_ragged(local, n_total, group=None):
    """All-gather per-unit results of a contiguous shard into the full array (every rank gets it).

    `local` is a 1-D or 2-D torch tensor holding this rank's shard (rows = units).  Shards may differ in
    size by one unit, so they are padded to the largest shard for the collective and trimmed afterwards.
    """
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    sizes = shard_sizes(n_total, world)
    assert local.shape[0] == sizes[dist.get_rank(group)], "shard size does not match shard_range()"
    mx = max(sizes)
    pad_shape = (mx,) + tuple(local.shape[1:])
    buf = torch.zeros(pad_shape, dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    parts = [out[r * mx: r * mx + sizes[r]] for r in range(world)]
    return torch.cat(parts, dim=0)


def merge_topk(dist_parts, idx_parts, k):
    """k-way merge of per-rank candidates with the reference tie rule.

    dist_parts / idx_parts: lists (one per rank) of arrays [nqueries][k_r]; idx are GLOBAL node indices,
    -1 marks padding.  Returns (idx [nqueries][k] int64, dist [nqueries][k] float64), nearest first,
    ordered by the key (distance, node index); padded with -1 / +inf when fewer than k candidates exist.
    """
    d = np.concatenate([np.asarray(x, dtype=np.float64) for x in dist_parts], axis=1)
    i = np.concatenate([np.asarray(x, dtype=np.int64) for x in idx_parts], axis=1)
    d = np.where(i < 0, np.inf, d)
    big = np.iinfo(np.int64).max
    ikey = np.where(i < 0, big, i)
    order = np.lexsort((ikey, d), axis=1)[:, :k]  # primary distance, secondary node index
    out_i = np.take_along_axis(i, order, axis=1)
    out_d = np.take_along_axis(d, order, axis=1)
    if out_i.shape[1] < k:
        pad = k - out_i.shape[1]
        out_i = np.concatenate([out_i, -np.ones((out_i.shape[0], pad), dtype=np.int64)], axis=1)
        out_d = np.concatenate([out_d, np.full((out_d.shape[0], pad), np.inf)], axis=1)
    out_d = np.where(out_i < 0, np.inf, out_d)
    return out_i, out_d


def all_gather_topk(local_idx, local_dist, node_offset, k, group=None):
    """Gather every rank's local top-k (local node indices + distances) and merge.

    local_idx [nqueries][k] int64 are indices into THIS rank's node shard (-1 padded); `node_offset` is
    the global index of the shard's first node.  Returns the global (idx, dist) on every rank.
    """
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    li = torch.as_tensor(local_idx, dtype=torch.int64).clone()
    li[li >= 0] += int(node_offset)
    ld = torch.as_tensor(local_dist, dtype=torch.float64).clone()
    nq = li.shape[0]
    gi = torch.empty((world * nq,) + tuple(li.shape[1:]), dtype=torch.int64, device=li.device)
    gd = torch.empty((world * nq,) + tuple(ld.shape[1:]), dtype=torch.float64, device=ld.device)
    dist.all_gather_into_tensor(gi, li.contiguous(), group=group)  # rank-major concatenation along dim 0
    dist.all_gather_into_tensor(gd, ld.contiguous(), group=group)
    if gi.is_cuda:  # merge on the device (knnMergeShardsKernel): results stay in HBM
        return merge_topk_device(gi.view(world, nq, -1), gd.view(world, nq, -1), k)
    gi = gi.view((world, nq) + tuple(li.shape[1:])).cpu().numpy()
    gd = gd.view((world, nq) + tuple(ld.shape[1:])).cpu().numpy()
    return merge_topk([gd[r] for r in range(world)], [gi[r] for r in range(world)], k)


def merge_topk_device(gi, gd, k):
    """gi / gd: CUDA tensors [world][nqueries][k] (global indices, -1 padded; distances), every list nearest first.
    Returns CUDA tensors (idx [nqueries][k] int64, dist [nqueries][k] float64): hppb200_knn_merge_shards on torch's
    current stream."""
    import ctypes as C
    import torch
    from . import lib, _check
    world, nq, kk = gi.shape
    assert kk == k and gi.is_contiguous() and gd.is_contiguous()
    out_i = torch.empty((nq, k), dtype=torch.int64, device=gi.device)
    out_d = torch.empty((nq, k), dtype=torch.float64, device=gi.device)
    stream = torch.cuda.current_stream(gi.device).cuda_stream
    _check(lib().hppb200_knn_merge_shards(C.c_void_p(stream), C.c_void_p(gd.data_ptr()), C.c_void_p(gi.data_ptr()), C.c_int(world),
                                          C.c_size_t(nq), C.c_int(k), C.c_void_p(out_i.data_ptr()), C.c_void_p(out_d.data_ptr())))
    return out_i, out_d


def all_gather_mask_packed(valid, group=None, unpack=True, bits=None, allbits=None):
    """valid: CUDA uint8 tensor [n] (same n on every rank).  Bit-packs it (hppb200_pack_mask), all-gathers the words
    (n / 8 bytes per rank instead of n) and, if `unpack`, expands them to uint8 [world * n]; all on torch's current stream.
    Returns (mask or None, allbits int32 [world * ceil(n / 32)]); rank r's words start at r * ceil(n / 32).
    `bits` / `allbits`: optional preallocated int32 buffers (a timed loop should not allocate)."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from . import lib, _check
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    n = valid.shape[0]
    words = (n + 31) // 32
    if bits is None:
        bits = torch.empty(words, dtype=torch.int32, device=valid.device)
    stream = torch.cuda.current_stream(valid.device).cuda_stream
    _check(lib().hppb200_pack_mask(C.c_void_p(stream), C.c_void_p(valid.data_ptr()), C.c_size_t(n), C.c_void_p(bits.data_ptr())))
    if world == 1:
        allbits = bits
    else:
        if allbits is None:
            allbits = torch.empty(world * words, dtype=torch.int32, device=valid.device)
        dist.all_gather_into_tensor(allbits, bits, group=group)
    if not unpack:
        return None, allbits
    return unpack_mask(allbits, n, world), allbits


def unpack_mask(allbits, n, world):
    """inverse of the packing above: uint8 [world * n]"""
    import ctypes as C
    import torch
    from . import lib, _check
    words = (n + 31) // 32
    out = torch.empty(world * n, dtype=torch.uint8, device=allbits.device)
    stream = torch.cuda.current_stream(allbits.device).cuda_stream
    for r in range(world):  # (n need not be a multiple of 32: every rank's words start at r * words)
        _check(lib().hppb200_unpack_mask(C.c_void_p(stream), C.c_void_p(allbits.data_ptr() + 4 * r * words), C.c_size_t(n),
                                         C.c_void_p(out.data_ptr() + r * n)))
    return out


class PeerMaskExchange:
    """hppb200_exchange_*: the mask gather as ONE kernel over NVLink peer memory (bit-pack + stores into every rank's buffer +
    epoch flags) instead of hppb200_pack_mask + an NCCL all-gather.  Ranks of one node; the IPC handles travel through
    torch.distributed.  `create` returns None on every rank when any rank cannot set it up (the caller keeps the NCCL form)."""

    class _Raw:
        def __init__(self, ptr, nwords):
            self.__cuda_array_interface__ = {"shape": (int(nwords),), "typestr": "<i4", "data": (int(ptr), False), "version": 2}

    def __init__(self, handle, rank, world):
        self.h, self.rank, self.world = handle, rank, world

    @classmethod
    def create(cls, device_index, max_units, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from . import lib
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        dev = torch.device("cuda", device_index)
        h = C.c_void_p()
        mine = (C.c_ubyte * 64)()
        ok = lib().hppb200_exchange_create(C.c_int(device_index), C.c_int(rank), C.c_int(world), C.c_size_t(max_units), C.byref(h), mine) == 0
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            if ok:
                lib().hppb200_exchange_destroy(h)
            return None
        hs = torch.empty(world * 64, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(hs, torch.tensor(list(mine), dtype=torch.uint8, device=dev), group=group)
        blob = bytes(hs.cpu().numpy().tobytes())
        ok = lib().hppb200_exchange_open_peers(h, C.c_char_p(blob)) == 0
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            lib().hppb200_exchange_destroy(h)
            return None
        return cls(h, rank, world)

    def gather(self, valid):
        """valid: CUDA uint8 [n] (same n on every rank), on torch's current stream.  Returns an int32 tensor VIEW of
        world * ceil(n / 32) words in the exchange buffer (valid until the second next call)."""
        import ctypes as C
        import torch
        from . import lib, _check
        n = valid.shape[0]
        out = C.c_void_p()
        stream = torch.cuda.current_stream(valid.device).cuda_stream
        _check(lib().hppb200_exchange_mask(self.h, C.c_void_p(stream), C.c_void_p(valid.data_ptr()), C.c_size_t(n), C.byref(out)))
        return torch.as_tensor(self._Raw(out.value, self.world * ((n + 31) // 32)), device=valid.device)

    def timed_out(self):
        """synchronises; True when a wait gave up (a peer never delivered)"""
        import ctypes as C
        from . import lib, _check
        t = C.c_int(0)
        _check(lib().hppb200_exchange_status(self.h, C.byref(t)))
        return t.value != 0

    def close(self):
        from . import lib
        if self.h:
            lib().hppb200_exchange_destroy(self.h)
            self.h = None
