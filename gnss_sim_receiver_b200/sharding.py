"""Multi-GPU layout of the two hot paths: one process per GPU, units sharded, no data-path collective.

Tracking: channels are independent (the first cross-channel consumer is Hybrid_Observables, downstream of the path), so a
receiver's channels are split over ranks — by antenna first (every IQ stream then lives on exactly one GPU), then channel
`c` of an antenna's list to rank `c mod G` among the ranks serving that antenna.  The reference's only multi-GPU logic is
`rand() % num_devices` per channel (src/algorithms/tracking/libs/cuda_multicorrelator.cu:136-154).

Acquisition: whole PRN grids stay on one rank (PRN `p` -> rank `p mod G`); the per-PRN result struct is all that leaves a
rank.  For the few-PRN / low-latency case a sweep can instead be tiled over Doppler bins; the tiles are merged with the
packed-key rule below, which preserves acquisition_core's first-index tie rule (pcps_acquisition.cc:496-597,600-697:
volk index_max over the grid rows in bin order, first maximum wins).

torch.distributed is plumbing only: barrier, max over ranks for timing, and a host gather of the small result records
(`gloo` on CPU in the tests, `nccl` on the GPUs).
"""
import numpy as np


# ---------------------------------------------------------------------------------------------------------- partitioning
def channel_shard(n_channels, rank, world, mode="mod"):
    """Indices of the channels rank `rank` tracks.  mode "mod": c -> c mod G;  "block": contiguous blocks of ceil(C/G)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world {world}")
    c = np.arange(int(n_channels), dtype=np.int64)
    if mode == "mod":
        return c[c % world == rank]
    if mode == "block":
        per = -(-int(n_channels) // world)
        return c[rank * per:(rank + 1) * per]
    raise ValueError(f"unknown mode {mode!r}")


def antenna_shard(channel_antenna, rank, world):
    """Channel indices for rank `rank` when channels belong to antennas (IQ streams): antennas are dealt to ranks first
    (antenna a -> rank a mod G) so each stream is uploaded to one GPU only; with fewer antennas than ranks the ranks
    sharing an antenna split its channels `mod` their count.  Returns (channels, antennas_this_rank_needs)."""
    channel_antenna = np.asarray(channel_antenna, dtype=np.int64)
    antennas = np.unique(channel_antenna)
    na = len(antennas)
    if na >= world:
        mine = antennas[np.arange(na) % world == rank]
        ch = np.nonzero(np.isin(channel_antenna, mine))[0]
        return ch.astype(np.int64), mine
    # fewer streams than GPUs: ranks r with r mod na == a serve antenna a
    a = antennas[rank % na]
    peers = [r for r in range(world) if r % na == rank % na]
    idx = np.nonzero(channel_antenna == a)[0]
    ch = idx[np.arange(len(idx)) % len(peers) == peers.index(rank)]
    return ch.astype(np.int64), np.array([a], dtype=np.int64)


def prn_shard(prns, rank, world):
    """PRNs (any iterable of ids) searched on rank `rank`: position p in the list -> rank p mod G."""
    prns = list(prns)
    return [p for i, p in enumerate(prns) if i % world == rank]


def doppler_tiles(n_bins, world):
    """Bin ranges [lo, hi) per rank for a Doppler-tiled sweep: bins [k*B/G, (k+1)*B/G)."""
    edges = [(k * int(n_bins)) // world for k in range(world + 1)]
    return [(edges[k], edges[k + 1]) for k in range(world)]


# ------------------------------------------------------------------------------------- Doppler-tile merge (packed key rule)
def pack_peak_key(peak, flat_index, total_cells):
    """64-bit key whose maximum is the grid maximum with the LOWEST flat index among equal peaks.
    peak: non-negative float32 (|.|^2 values order like their bit patterns); flat_index = bin * N_eff + indext."""
    bits = np.asarray(peak, dtype=np.float32).view(np.uint32).astype(np.uint64)
    if np.any(np.asarray(peak) < 0):
        raise ValueError("magnitudes are non-negative")
    tail = np.uint64(total_cells) - np.asarray(flat_index, dtype=np.uint64)
    if np.any(tail >= np.uint64(1) << np.uint64(32)) or np.any(np.asarray(flat_index) >= total_cells):
        raise ValueError("grid too large for the 32-bit index field")
    return (bits << np.uint64(32)) | tail


def unpack_peak_key(key, total_cells):
    key = np.asarray(key, dtype=np.uint64)
    peak = (key >> np.uint64(32)).astype(np.uint32).view(np.float32)
    flat = np.uint64(total_cells) - (key & np.uint64(0xFFFFFFFF))
    return peak, flat.astype(np.int64)


def tile_key(grid_tile, bin_lo, n_bins, neff):
    """Packed key of one rank's tile (rows = its Doppler bins, columns = the first N_eff code phases)."""
    g = np.asarray(grid_tile, dtype=np.float32)[:, :neff]
    if g.size == 0:
        return np.uint64(0)
    j = int(np.argmax(g))  # first maximum in row-major order == bin order then code-phase order
    b, i = divmod(j, g.shape[1])
    return pack_peak_key(g[b, i], (bin_lo + b) * neff + i, n_bins * neff)


def merge_doppler_tiles(dist, grid_tile, bin_lo, n_bins, neff, num_dwells=1, device="cpu"):
    """CFAR statistic (max_to_input_power_statistic, pcps_acquisition.cc:496-538) of a sweep whose grid rows are spread
    over the ranks.  Exchange: one MAX all-reduce of the 64-bit key, then one SUM all-reduce of the opposite bin's power
    (8 bytes each).  Returns dict(indext, doppler_index, peak, input_power, test_statistics) on every rank."""
    import torch
    key = int(tile_key(grid_tile, bin_lo, n_bins, neff))
    t = torch.tensor([key], dtype=torch.int64, device=device)  # top bit is the float sign: always 0, fits int64
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    peak, flat = unpack_peak_key(np.array([t.item()], dtype=np.uint64), n_bins * neff)
    doppler_index, indext = divmod(int(flat[0]), neff)
    # input power comes from the bin opposite to the peak's (:515-527): whoever owns that row contributes its sum
    opp = (doppler_index + n_bins // 2) % n_bins
    g = np.asarray(grid_tile, dtype=np.float32)
    local = 0.0
    if bin_lo <= opp < bin_lo + g.shape[0]:
        local = float(np.sum(g[opp - bin_lo, :neff], dtype=np.float32))
    p = torch.tensor([local], dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(p, op=dist.ReduceOp.SUM)
    power = np.float32(p.item() / neff / 2.0 / num_dwells)  # :519-521: float accumulate, then / N_eff / 2 / dwells
    stat = np.float32(peak[0]) / power if power > 0 else np.float32(0)
    return dict(indext=indext, doppler_index=doppler_index, peak=float(peak[0]), input_power=float(power), test_statistics=float(stat))


# ----------------------------------------------------------------- Doppler-tiled sweep on the GPUs: merge of per-bin cell reductions
def merge_cells(vmax, imax, sums, neff, num_dwells=1, doppler_base=0, doppler_step=0, samples_per_code=None):
    """Statistic of a sweep from its per-(slot, bin) reductions (gb200_acq_read_cells), bins in sweep order — the arithmetic of
    acq_stats_kernel / max_to_input_power_statistic (pcps_acquisition.cc:496-528), operation for operation in fp32 / float64:
    first Doppler bin wins among equal maxima (strict '>' over bins in order), first code phase within a bin (already in imax),
    input power = sum of the OPPOSITE bin / N_eff / 2 / dwells.  Returns a list of dicts, one per slot."""
    vmax = np.asarray(vmax, dtype=np.float32)
    imax = np.asarray(imax, dtype=np.int64)
    sums = np.asarray(sums, dtype=np.float32)
    n_slots, n_bins = vmax.shape
    rows = np.arange(n_slots)
    b = np.argmax(vmax, axis=1)  # first maximum in bin order
    best = vmax[rows, b]
    idx = imax[rows, b]
    dead = ~(best > 0)           # an all-zero grid leaves (0, 0): the reference starts from grid_maximum = 0 with a strict '>'
    best = np.where(dead, np.float32(0), best).astype(np.float32)
    b = np.where(dead, 0, b)
    idx = np.where(dead, 0, idx)
    opp = (b + n_bins // 2) % n_bins
    power = ((sums[rows, opp] / np.float32(neff)).astype(np.float64) / 2.0 / num_dwells).astype(np.float32)  # float / int, then double (:519)
    with np.errstate(divide="ignore", invalid="ignore"):
        stat = (best / power).astype(np.float32)
    delay = np.fmod(idx.astype(np.float32), np.float32(samples_per_code)) if samples_per_code is not None else None
    out = []
    for s in range(n_slots):
        r = dict(indext=int(idx[s]), doppler_index=int(b[s]), doppler_hz=int(doppler_base + doppler_step * int(b[s])), peak=float(best[s]), input_power=float(power[s]),
                 test_statistics=float(stat[s]))
        if delay is not None:
            r["acq_delay_samples"] = float(delay[s])
        out.append(r)
    return out


def gather_cell_tiles(dist, vmax, imax, sums, world, device="cpu", widths=None):
    """All-gather of every rank's Doppler tile of cell reductions ([n_slots, bins_of_this_rank] each; tiles may differ in width) into
    the full [n_slots, n_bins] tables on EVERY rank, tiles in rank order = bin order (doppler_tiles).  One collective of
    3 x 4 bytes per (slot, bin): NCCL on the GPUs, gloo in the CPU tests."""
    if dist is None or world == 1:
        return np.asarray(vmax), np.asarray(imax), np.asarray(sums)
    import torch
    n_slots, nb = np.asarray(vmax).shape
    if widths is None:  # tile widths of the other ranks (known beforehand when the tiles come from doppler_tiles: pass them and save a collective)
        wt = torch.zeros(world, dtype=torch.int64, device=device)
        mine = torch.tensor([nb], dtype=torch.int64, device=device)
        dist.all_gather_into_tensor(wt, mine)
        widths = [int(w) for w in wt.tolist()]
    wmax = max(widths)
    pack = np.zeros((3, n_slots, wmax), np.float32)
    pack[0, :, :nb] = vmax
    pack[1, :, :nb] = np.ascontiguousarray(imax, dtype=np.int32).view(np.float32)  # the index bits travel in a float lane
    pack[2, :, :nb] = sums
    t = torch.from_numpy(pack).to(device).reshape(-1)
    allt = torch.empty(world * t.numel(), dtype=t.dtype, device=device)
    dist.all_gather_into_tensor(allt, t)
    a = allt.cpu().numpy().reshape((world,) + pack.shape)
    v = np.concatenate([a[r, 0, :, :widths[r]] for r in range(world)], axis=1)
    i = np.concatenate([np.ascontiguousarray(a[r, 1, :, :widths[r]]).view(np.int32) for r in range(world)], axis=1)
    sm = np.concatenate([a[r, 2, :, :widths[r]] for r in range(world)], axis=1)
    return v, i, sm


# --------------------------------------------------------------------------------------------------------- plumbing
def reduce_max(dist, x, device="cpu"):
    """Max over ranks of a host scalar (timing: the slowest rank defines the step)."""
    if dist is None:
        return float(x)
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_records(dist, records, rank, world, key=None):
    """Host gather of per-rank numpy record arrays to rank 0 (correlator outputs / Gnss_Synchro-like records / per-PRN
    acquisition results).  Returns the concatenation (sorted by `key` if given) on rank 0, None elsewhere."""
    if dist is None or world == 1:
        out = np.asarray(records)
    else:
        bucket = [None] * world if rank == 0 else None
        dist.gather_object(np.asarray(records), bucket, dst=0)
        if rank != 0:
            return None
        out = np.concatenate([b for b in bucket if b is not None and len(b)])
    if key is not None and len(out):
        out = out[np.argsort(out[key], kind="stable")]
    return out
