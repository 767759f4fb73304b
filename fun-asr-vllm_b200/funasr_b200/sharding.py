"""Utterance sharding across the GPUs of one box: no data-path collective.

``shard_contiguous`` is the reference's own scheme (``/root/reference/infer_kaldidata.py:493-544``: N contiguous parts
of wav.scp, the remainder spread over the first parts).  ``shard_lpt`` is the B200 default: longest-processing-time
first on the cost model of SURVEY.md §8e so variable-length sets stay balanced.  ``gather_results`` is the host-side
merge (``infer_kaldidata.py:1013-1062`` merges files sorted by utterance id)."""
from __future__ import annotations

import heapq


def shard_contiguous(n_items: int, world_size: int) -> list[list[int]]:
    base, rem = divmod(n_items, world_size)
    out, start = [], 0
    for r in range(world_size):
        n = base + (1 if r < rem else 0)
        out.append(list(range(start, start + n)))
        start += n
    return out


def utterance_cost(n_lfr: int) -> float:
    """GFLOP of encoder+adaptor for one utterance of T LFR frames: 0.4657 T (GEMM) + 1.516e-4 T^2 (attention)."""
    return 0.4657152 * n_lfr + 1.51552e-4 * n_lfr * n_lfr


def shard_lpt(n_lfr: list[int], world_size: int) -> list[list[int]]:
    """Greedy LPT: sort by cost descending, always give the next utterance to the least-loaded rank.  Deterministic
    (ties broken by index) so every rank computes the same partition without communicating."""
    order = sorted(range(len(n_lfr)), key=lambda i: (-utterance_cost(n_lfr[i]), i))
    heap = [(0.0, r) for r in range(world_size)]
    heapq.heapify(heap)
    out = [[] for _ in range(world_size)]
    for i in order:
        load, r = heapq.heappop(heap)
        out[r].append(i)
        heapq.heappush(heap, (load + utterance_cost(n_lfr[i]), r))
    for part in out:
        part.sort()
    return out


def gather_results(local: dict, group=None) -> dict:
    """Merge per-rank ``{utt_id: result}`` dicts on every rank, ordered by utterance id.  Host objects only (text, lengths);
    tensors never cross ranks."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return dict(sorted(local.items()))
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, local, group=group)
    merged = {}
    for p in parts:
        merged.update(p)
    return dict(sorted(merged.items()))


def bind_to_gpu_numa(device_index: int) -> list[int] | None:
    """One process per GPU (``/root/reference/infer_kaldidata.py:1116-1122``): pin this process to the CPU cores NVML reports as
    local to GPU ``device_index`` (``CUDA_VISIBLE_DEVICES`` order is honoured), so that the pinned staging buffers it allocates
    next land on the NUMA node of that GPU's PCIe root.  Best effort: returns the cores used, or ``None`` when NVML or the
    affinity call is unavailable or the intersection with the allowed cores is empty (nothing is changed then)."""
    import os

    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = device_index
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if device_index < len(ids) and ids[device_index].isdigit():
                phys = int(ids[device_index])
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        n_cpu = os.cpu_count() or 64
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        ideal = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        cores = sorted(ideal & os.sched_getaffinity(0))
        if not cores:
            return None
        os.sched_setaffinity(0, cores)
        return cores
    except Exception:
        return None
