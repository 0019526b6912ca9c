"""Multi-GPU plumbing of the path (one process per GPU, torch.distributed).

Extraction shards by frame with no collective at all (frames are independent: SURVEY.md §8e).  The only exchange
step in the path family is the cross-frame descriptor sweep (BASELINE.json configs[4]): every rank all-gathers the
fixed-capacity descriptor slabs + counts of its frames, then matches its share of the frame pairs with the Hamming
kernel.  torch is imported lazily: nothing else in the package depends on it."""
import numpy as np


def frames_for_rank(nframes: int, rank: int, world: int) -> np.ndarray:
    """Round-robin frame ownership: frame f belongs to rank f mod world."""
    return np.arange(rank, nframes, world, dtype=np.int64)


def global_frame_index(local_index: int, rank: int, world: int) -> int:
    return local_index * world + rank


def pair_partition(nframes: int, rank: int, world: int) -> np.ndarray:
    """Upper-triangle frame pairs (i < j) dealt block-cyclically to the ranks; returns this rank's (k, 2) array."""
    i, j = np.triu_indices(nframes, k=1)
    sel = (np.arange(len(i)) % world) == rank
    return np.stack([i[sel], j[sel]], 1).astype(np.int64)


def allgather_descriptors(desc, counts, group=None):
    """desc: (F_local, cap, 32) uint8 tensor, counts: (F_local,) int32 tensor (same F_local and cap on every rank).
    Returns (desc_all (F_local*world, cap, 32), counts_all) with GLOBAL frame order (frame g = local*world + rank)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    F, cap = desc.shape[0], desc.shape[1]
    d_all = torch.empty((world,) + tuple(desc.shape), dtype=desc.dtype, device=desc.device)
    c_all = torch.empty((world, F), dtype=counts.dtype, device=counts.device)
    dist.all_gather_into_tensor(d_all.view(-1), desc.contiguous().view(-1), group=group)
    dist.all_gather_into_tensor(c_all.view(-1), counts.contiguous().view(-1), group=group)
    # rank-major -> global frame order
    d_all = d_all.permute(1, 0, 2, 3).reshape(F * world, cap, 32).contiguous()
    c_all = c_all.permute(1, 0).reshape(F * world).contiguous()
    return d_all, c_all


def sweep_pairs(desc_all, counts_all, pairs, device_index, th_low=50, nnratio=0.9):
    """Matches every (i, j) of `pairs` in ONE launch of the batched Hamming kernel (device tensors); returns the number
    of accepted matches per pair under ORBmatcher's test best <= TH_LOW and best < nnratio * second
    (src/ORBmatcher.cc:454-456)."""
    import torch
    from . import matcher
    dev = desc_all.device
    cap = desc_all.shape[1]
    p = torch.as_tensor(np.ascontiguousarray(pairs, np.int32)).to(dev)
    acc = torch.zeros(max(len(pairs), 1), dtype=torch.int32, device=dev)
    c32 = counts_all.to(torch.int32).contiguous()
    stream = torch.cuda.current_stream(dev).cuda_stream
    matcher.hamming_knn2_pairs_raw(desc_all.data_ptr(), c32.data_ptr(), cap, p.data_ptr(), len(pairs), None, None, None, acc.data_ptr(),
                                   nnratio, th_low, device=device_index, stream=stream)
    return acc[:len(pairs)]
