"""SURVEY 8e on real GPUs (needs >= 2 devices; skipped on a one-GPU box): ONE acquisition sweep split over two ranks

  * by PRN (slot p -> rank p mod G, whole grids stay local, one result struct per PRN gathered), and
  * by Doppler tile (rank k sweeps bins [kB/G, (k+1)B/G) of every PRN through gb200_acq_set_doppler_bins; the per-(PRN, bin) cell
    reductions — maximum, first index, row sum: 12 bytes each — are all-gathered over NCCL and merged with the reference's first-bin /
    first-index rule, sharding.merge_cells),

and the merged result must equal the single-GPU sweep BIT FOR BIT (index, bin, peak, input power, statistic): sharding adds no arithmetic.
One process per GPU, torch.distributed (nccl) as plumbing, rendezvous on 127.0.0.1.
"""
import os
import socket
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FS, N, DMAX, DSTEP, PRNS = 25e6, 25000, 10000, 250, 16


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _scene():
    import bench
    sig, sats = bench.acq_signal(N, 41)
    codes = bench.gps_codes()
    sampled = [bench.gps_code_sampled(codes[p], FS, N) for p in range(PRNS)]
    return sig, sampled


def _worker(rank, world, port, outdir):
    import torch
    import torch.distributed as dist
    from gnss_sim_receiver_b200 import sharding as S
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        sig, sampled = _scene()
        spchip = int(np.ceil(FS / 1.023e6))
        kw = dict(fs_in=FS, fft_size=N, samples_per_code=float(N), samples_per_chip=spchip, device=rank)
        # ---- single-GPU reference sweep (rank 0)
        single = None
        if rank == 0:
            e = AcquisitionEngine(max_prns=PRNS, **kw)
            e.set_doppler_grid(DMAX, DSTEP)
            for p in range(PRNS):
                e.set_local_code(p, sampled[p])
            r = e.run(sig)
            single = np.array([[x.indext, x.doppler_index, x.doppler_hz, np.float32(x.peak).view(np.uint32), np.float32(x.input_power).view(np.uint32),
                                np.float32(x.test_statistics).view(np.uint32), np.float32(x.acq_delay_samples).view(np.uint32)] for x in r], dtype=np.int64)
            bins = e.num_doppler_bins
            e.close()
        else:
            bins = int(np.ceil(2 * DMAX / DSTEP))
        # ---- by PRN
        mine = [p for p in range(PRNS) if p % world == rank]
        e1 = AcquisitionEngine(max_prns=len(mine), **kw)
        e1.set_doppler_grid(DMAX, DSTEP)
        for i, p in enumerate(mine):
            e1.set_local_code(i, sampled[p])
        r = e1.run(sig)
        rec = torch.tensor([[p, x.indext, x.doppler_index, x.doppler_hz, int(np.float32(x.peak).view(np.uint32)), int(np.float32(x.input_power).view(np.uint32)),
                             int(np.float32(x.test_statistics).view(np.uint32)), int(np.float32(x.acq_delay_samples).view(np.uint32))] for p, x in zip(mine, r)],
                           dtype=torch.int64, device=f"cuda:{rank}")
        e1.close()
        allr = [torch.empty_like(rec) for _ in range(world)]
        dist.all_gather(allr, rec)
        by_prn = np.concatenate([t.cpu().numpy() for t in allr])
        by_prn = by_prn[np.argsort(by_prn[:, 0])][:, 1:]
        # ---- by Doppler tile
        lo, hi = S.doppler_tiles(bins, world)[rank]
        e2 = AcquisitionEngine(max_prns=PRNS, **kw)
        e2.set_doppler_bins([float(-DMAX + DSTEP * k) for k in range(lo, hi)])
        for p in range(PRNS):
            e2.set_local_code(p, sampled[p])
        e2.run(sig)
        v, im, sm = e2.read_cells(PRNS)
        e2.close()
        v, im, sm = S.gather_cell_tiles(dist, v, im, sm, world, device=f"cuda:{rank}")
        merged = S.merge_cells(v, im, sm, N, 1, -DMAX, DSTEP, float(N))
        tiled = np.array([[m["indext"], m["doppler_index"], m["doppler_hz"], np.float32(m["peak"]).view(np.uint32), np.float32(m["input_power"]).view(np.uint32),
                           np.float32(m["test_statistics"]).view(np.uint32), np.float32(m["acq_delay_samples"]).view(np.uint32)] for m in merged], dtype=np.int64)
        dist.barrier()
        np.save(os.path.join(outdir, f"tiled{rank}.npy"), tiled)
        if rank == 0:
            np.save(os.path.join(outdir, "single.npy"), single)
            np.save(os.path.join(outdir, "by_prn.npy"), by_prn)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs")
@pytest.mark.timeout(600)
def test_one_sweep_over_two_gpus_equals_the_single_gpu_sweep_bit_for_bit():
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    with tempfile.TemporaryDirectory() as d:
        ctx = mp.get_context("spawn")
        procs = [ctx.Process(target=_worker, args=(r, world, port, d)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(500)
        for p in procs:
            if p.is_alive():
                p.kill()  # the exact processes started above
                pytest.fail("rank did not finish")
            assert p.exitcode == 0
        single, by_prn = np.load(os.path.join(d, "single.npy")), np.load(os.path.join(d, "by_prn.npy"))
        tiled = [np.load(os.path.join(d, f"tiled{r}.npy")) for r in range(world)]
    assert np.array_equal(by_prn, single)       # PRN-sharded: same units, only their placement changed
    for t in tiled:                             # Doppler-tiled: every rank ends with the single-GPU answer, bit for bit
        assert np.array_equal(t, single), np.nonzero(t != single)
