"""SURVEY 8e on CPU: the N>1 host logic (channel / antenna / PRN sharding, Doppler-tile merge, record gather, max-over-ranks
timing) run as world_size-2 `gloo` process groups.  The per-unit compute on each rank is the oracle (the checker stands in for
the GPU here — the sharding layer never looks inside a unit), and rank 0's gathered result must equal the single-process one
bit for bit: sharding adds no arithmetic.
"""
import os
import socket
import sys
import tempfile

import numpy as np
import pytest

import _oracle as O
from gnss_sim_receiver_b200 import sharding as S

FS = 4e6
N = 4000


# ------------------------------------------------------------------------------------------------ pure partitioning rules
def test_channel_shard_is_a_partition():
    for world in (1, 2, 3, 8):
        for mode in ("mod", "block"):
            got = np.concatenate([S.channel_shard(37, r, world, mode) for r in range(world)])
            assert sorted(got.tolist()) == list(range(37))
    assert S.channel_shard(8, 1, 2).tolist() == [1, 3, 5, 7]
    assert S.channel_shard(8, 1, 2, "block").tolist() == [4, 5, 6, 7]
    assert S.channel_shard(0, 0, 2).tolist() == []  # empty receiver
    assert S.channel_shard(1, 1, 2).tolist() == []  # fewer channels than ranks
    with pytest.raises(ValueError):
        S.channel_shard(4, 2, 2)


def test_antenna_shard_keeps_a_stream_on_one_gpu():
    # C5 shape in small: 8 antennas x 5 channels over 4 ranks -> 2 antennas per rank, every stream on exactly one rank
    ant = np.repeat(np.arange(8), 5)
    seen = {}
    allch = []
    for r in range(4):
        ch, mine = S.antenna_shard(ant, r, 4)
        assert len(mine) == 2 and set(ant[ch]) == set(mine.tolist())
        for a in mine:
            assert a not in seen
            seen[a] = r
        allch += ch.tolist()
    assert sorted(allch) == list(range(40))
    # fewer antennas than ranks: the ranks sharing a stream split its channels
    ant = np.array([0] * 6 + [1] * 3)
    allch = []
    for r in range(4):
        ch, mine = S.antenna_shard(ant, r, 4)
        assert len(mine) == 1 and np.all(ant[ch] == mine[0])
        allch += ch.tolist()
    assert sorted(allch) == list(range(9))


def test_merge_cells_tie_rules_and_statistic():
    """merge_cells = acq_stats_kernel's arithmetic on the host: first bin among equal maxima, the opposite bin's sum as input power"""
    v = np.zeros((2, 8), np.float32)
    i = np.zeros((2, 8), np.int32)
    s = np.ones((2, 8), np.float32) * 100
    v[0, 2] = v[0, 5] = 7.0
    i[0, 2], i[0, 5] = 40, 3
    s[0, 6] = 50.0   # opposite of bin 2 in an 8-bin grid
    r = S.merge_cells(v, i, s, neff=10, num_dwells=1, doppler_base=-1000, doppler_step=250, samples_per_code=16.0)
    assert (r[0]["indext"], r[0]["doppler_index"], r[0]["doppler_hz"]) == (40, 2, -500)
    assert r[0]["input_power"] == np.float32(np.float64(np.float32(50.0) / np.float32(10)) / 2.0) and r[0]["test_statistics"] == np.float32(7.0) / np.float32(2.5)
    assert r[0]["acq_delay_samples"] == 8.0
    # an all-zero grid leaves (0, 0) like the reference's strict '>' from grid_maximum = 0 (pcps_acquisition.cc:505-514)
    assert (r[1]["indext"], r[1]["doppler_index"], r[1]["peak"]) == (0, 0, 0.0)


def test_prn_shard_and_doppler_tiles():
    prns = list(range(1, 33))
    parts = [S.prn_shard(prns, r, 3) for r in range(3)]
    assert sorted(sum(parts, [])) == prns and parts[0][:3] == [1, 4, 7]
    for world in (1, 2, 3, 7):
        t = S.doppler_tiles(80, world)
        assert t[0][0] == 0 and t[-1][1] == 80 and all(t[k][1] == t[k + 1][0] for k in range(world - 1))
    assert S.doppler_tiles(3, 4) == [(0, 0), (0, 1), (1, 2), (2, 3)]  # fewer bins than ranks: an empty tile


def test_packed_key_orders_like_first_index_max():
    rng = np.random.default_rng(0)
    g = rng.random((7, 50)).astype(np.float32)
    g[2, 9] = g[5, 1] = g[2, 30] = 2.0  # three equal maxima: (2, 9) is first in bin order
    keys = S.pack_peak_key(g.ravel(), np.arange(g.size), g.size)
    peak, flat = S.unpack_peak_key(keys.max(), g.size)
    assert peak == np.float32(2.0) and divmod(int(flat), 50) == (2, 9)
    assert int(S.tile_key(g, 0, 7, 50)) == int(keys.max())
    with pytest.raises(ValueError):
        S.pack_peak_key(np.float32(-1.0), 0, 10)
    with pytest.raises(ValueError):
        S.pack_peak_key(np.float32(1.0), 10, 10)


# ------------------------------------------------------------------------------------------ world_size-2 gloo process groups
def _scene():
    rng = np.random.default_rng(42)
    prns = [3, 11, 19, 27]
    codes = np.stack([O.gps_l1_ca_code(p) for p in prns])
    # two of the four PRNs are in the air
    x = O.gen_if(3 * N, 0, FS, codes[:2], [1250.0, -3000.0], [200.3, 811.9], [0.3, 2.1], [48.0, 45.0], seed=5)
    nch = 6
    trk = dict(ch_prn=rng.integers(0, 4, nch), off=rng.integers(0, N, nch), rem_carr=rng.uniform(0, 6.2, nch).astype(np.float32),
               phase_step=rng.uniform(-0.02, 0.02, nch).astype(np.float32), rem_code=rng.uniform(0, 1, nch).astype(np.float32),
               code_step=(1.023e6 / FS * (1 + rng.uniform(-3e-6, 3e-6, nch))).astype(np.float32))
    return prns, codes, x, trk


REC = np.dtype([("channel", np.int32), ("taps", np.complex64, (3,))])
ACQ = np.dtype([("prn", np.int32), ("indext", np.uint32), ("doppler_index", np.int32), ("stat", np.float32), ("delay", np.float32), ("doppler", np.float32)])
SHIFTS = np.array([-0.5, 0.0, 0.5], dtype=np.float32)


def _track(codes, x, trk, channels):
    out = np.zeros(len(channels), dtype=REC)
    for j, c in enumerate(channels):
        out["channel"][j] = c
        out["taps"][j] = O.mc_real_codes(x[trk["off"][c]:trk["off"][c] + N], codes[trk["ch_prn"][c]], SHIFTS, trk["rem_carr"][c], trk["phase_step"][c],
                                         trk["rem_code"][c], trk["code_step"][c])
    return out


def _acq_conf():
    return O.acq_conf(FS, N, N, 5000, 250, float(np.float32(FS) * np.float32(0.001)), 4)


def _acquire(x, prns):
    conf = _acq_conf()
    out = np.zeros(len(prns), dtype=ACQ)
    for j, p in enumerate(prns):
        fc = O.acq_set_local_code(conf, O.gps_l1_ca_code_sampled(p, int(FS), N))
        r = O.acq_statistic(conf, O.acq_dwell(conf, x[:N], fc))
        out[j] = (p, r.indext, r.doppler_index, r.test_statistics, r.acq_delay_samples, r.acq_doppler_hz)
    return out


def _worker(rank, world, port, outdir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        prns, codes, x, trk = _scene()
        # tracking: channels sharded, records gathered on rank 0
        mine = S.channel_shard(len(trk["off"]), rank, world)
        recs = S.gather_records(dist, _track(codes, x, trk, mine), rank, world, key="channel")
        # acquisition: PRNs sharded, one result record per PRN gathered
        acq = S.gather_records(dist, _acquire(x, S.prn_shard(prns, rank, world)), rank, world, key="prn")
        # acquisition, Doppler-tiled: one PRN, grid rows split over the ranks, packed-key merge
        conf = _acq_conf()
        fc = O.acq_set_local_code(conf, O.gps_l1_ca_code_sampled(prns[1], int(FS), N))
        grid = O.acq_dwell(conf, x[:N], fc)
        lo, hi = S.doppler_tiles(conf.num_doppler_bins, world)[rank]
        tiled = S.merge_doppler_tiles(dist, grid[lo:hi], lo, conf.num_doppler_bins, N)
        # tie across tiles: equal maxima in both halves -> the lower bin wins on every rank
        tie = np.zeros((conf.num_doppler_bins, N), dtype=np.float32)
        tie[3, 77] = tie[30, 5] = 9.0
        tied = S.merge_doppler_tiles(dist, tie[lo:hi], lo, conf.num_doppler_bins, N)
        # the form the GPUs use (bench.py --gpus N): every rank publishes (max, first index, sum) per (PRN, bin) of ITS tile
        # (gb200_acq_read_cells), one all-gather, then the reference's first-bin / first-index rule on the full table
        tile = grid[lo:hi]
        v, im, sm = tile.max(axis=1)[None, :], tile.argmax(axis=1).astype(np.int32)[None, :], tile.sum(axis=1, dtype=np.float32)[None, :]
        fv, fi, fs_ = S.gather_cell_tiles(dist, v, im, sm, world)
        cells = S.merge_cells(fv, fi, fs_, N, 1, -conf.doppler_max, conf.doppler_step, conf.samples_per_code)[0]
        np.save(os.path.join(outdir, f"cells{rank}.npy"), np.array([cells["indext"], cells["doppler_index"], cells["doppler_hz"], cells["peak"], cells["acq_delay_samples"]]))
        # timing plumbing: the slowest rank defines the step
        slow = S.reduce_max(dist, 1.0 + rank)
        dist.barrier()
        np.save(os.path.join(outdir, f"tiled{rank}.npy"), np.array([tiled["indext"], tiled["doppler_index"], tiled["test_statistics"], tied["indext"], tied["doppler_index"], slow]))
        if rank == 0:
            np.save(os.path.join(outdir, "trk.npy"), recs)
            np.save(os.path.join(outdir, "acq.npy"), acq)
        else:
            assert recs is None and acq is None
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(300)
def test_world_size_2_gloo_matches_single_process():
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    with tempfile.TemporaryDirectory() as d:
        ctx = mp.get_context("spawn")
        procs = [ctx.Process(target=_worker, args=(r, world, port, d)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(240)
        for p in procs:
            if p.is_alive():
                p.kill()  # the exact processes started above
                pytest.fail("rank did not finish")
            assert p.exitcode == 0
        trk, acq = np.load(os.path.join(d, "trk.npy")), np.load(os.path.join(d, "acq.npy"))
        tiled = [np.load(os.path.join(d, f"tiled{r}.npy")) for r in range(world)]
        cells = [np.load(os.path.join(d, f"cells{r}.npy")) for r in range(world)]
    prns, codes, x, t = _scene()
    one_trk = _track(codes, x, t, np.arange(len(t["off"])))
    one_acq = _acquire(x, prns)
    assert trk.tobytes() == one_trk.tobytes()  # bit-exact: same units, only their placement changed
    assert acq.tobytes() == one_acq.tobytes()
    # the two PRNs in the air are found where they were put (code start = delay / f_chip * fs)
    for j, (dly, dop) in enumerate([(200.3, 1250.0), (811.9, -3000.0)]):
        assert abs(acq["delay"][j] - dly / 1.023e6 * FS) <= 2 and abs(acq["doppler"][j] - dop) <= 125
    conf = _acq_conf()
    fc = O.acq_set_local_code(conf, O.gps_l1_ca_code_sampled(prns[1], int(FS), N))
    r = O.acq_statistic(conf, O.acq_dwell(conf, x[:N], fc))
    for v in tiled:  # every rank ends with the same merged answer
        assert (int(v[0]), int(v[1])) == (r.indext, r.doppler_index)
        assert abs(v[2] - r.test_statistics) <= 1e-5 * r.test_statistics
        assert (int(v[3]), int(v[4])) == (77, 3)
        assert v[5] == 2.0
    for c in cells:  # the all-gathered cell tables give the oracle's answer on every rank, peak value bit for bit
        assert (int(c[0]), int(c[1]), int(c[2])) == (r.indext, r.doppler_index, r.doppler_hz)
        assert np.float32(c[3]) == np.float32(r.peak) and np.float32(c[4]) == np.float32(r.acq_delay_samples)


def test_bench_reference_arm_runs_on_rank_0_only():
    """`bench.py --impl reference` under torchrun: ranks other than 0 exit 0 without work or output."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()))
    p = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       env=env, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""
