"""N > 1 host logic on CPU: two `gloo` ranks each render their interleaved-band shard (with the oracle standing in for
the GPU call) and rank 0 assembles the frame; also the frame round-robin of animations."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist

    import oracle_binding as ob
    from fractal_renderer_b200 import configs, sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = configs.c1().with_(width=160, height=107)  # 107 rows: ragged last band

        def render(c, shard):
            return ob.render_rows(c, rows=shard.image_rows(c.height), threads=1)["rgb"]

        frame = sharding.render_distributed(cfg, render, dist)
        frames = sharding.frame_of_rank(10, rank, world)
        if rank == 0:
            ref = ob.render_rows(cfg, threads=1)["rgb"]
            q.put(("frame_ok", bool(np.array_equal(frame, ref))))
        else:
            assert frame is None
        q.put(("frames", rank, frames))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_assemble():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(3)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ("frame_ok", True) in got
    fr = {g[1]: g[2] for g in got if g[0] == "frames"}
    assert fr[0] == [0, 2, 4, 6, 8] and fr[1] == [1, 3, 5, 7, 9]


def _shared_worker(rank, world, name, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    import fractal_renderer_b200 as frb
    from fractal_renderer_b200 import configs, sharding

    cfg = configs.c1().with_(width=160, height=107)
    n = cfg.width * cfg.height * 3

    def cpu_render_bands(c, shard, frame, fmt, device):  # stands in for frb_render_bands (there is no GPU here)
        rows = shard.image_rows(c.height)
        frame[:n].reshape(c.height, c.width, 3)[rows] = ob.render_rows(c, rows=rows, threads=1)["rgb"]

    with frb.SharedFrame(name, n + 8, world, rank) as sf:
        for rep in range(3):  # the barrier is reusable (sense reversing)
            img = sharding.render_into_shared_frame(cfg, sf, rank, world, cpu_render_bands)
            if rank == 0:
                ok = bool(np.array_equal(img, ob.render_rows(cfg, threads=1)["rgb"]))
                q.put(("shared_ok", rep, ok))
            sf.barrier()               # nobody overwrites the frame while rank 0 is still comparing
            sf.array[:n] = 0 if rank == 0 else sf.array[:n]
            sf.barrier()
        # a counter in the frame's spare bytes: every rank adds its number between barriers, 200 rounds
        cnt = sf.array[n:n + 8].view(np.uint64)
        for i in range(200):
            if i % world == rank:
                cnt[0] += 1
            sf.barrier()
        q.put(("count", rank, int(cnt[0])))
        sf.barrier()


def test_shared_frame_two_processes_assemble_one_image():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"/frb-test-{os.getpid()}"
    procs = [ctx.Process(target=_shared_worker, args=(r, world, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(3 + world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(g[2] for g in got if g[0] == "shared_ok") and sum(g[0] == "shared_ok" for g in got) == 3
    assert all(g[2] == 200 for g in got if g[0] == "count")
    assert not os.path.exists("/dev/shm" + name)


def test_shared_frame_left_over_from_a_crashed_job_is_detected():
    """A rank that attaches to the left-over frame of a crashed job (same name, magic still set) before this job's rank 0 has
    replaced it must not sit in the barrier for a minute, let alone render into memory nobody reads: the barrier notices that
    the file it mapped was removed and says so; a second open attaches to the new frame."""
    import time
    import fractal_renderer_b200 as frb

    name = f"/frb-stale-{os.getpid()}"
    crashed = frb.SharedFrame(name, 4096, 2, 0)           # the "crashed" job's rank 0: never closed, its file stays
    late = frb.SharedFrame(name, 4096, 2, 1)              # this job's rank 1 finds the old file first
    fresh = frb.SharedFrame(name, 4096, 2, 0)             # this job's rank 0 removes it and creates the frame anew
    try:
        t0 = time.perf_counter()
        with pytest.raises(frb.FrbError) as e:
            late.barrier()
        assert time.perf_counter() - t0 < 20.0
        assert "removed or replaced" in str(e.value)
        late.close()
        late = frb.SharedFrame(name, 4096, 2, 1)          # attaches to the new file
        fresh.array[:4] = (1, 2, 3, 4)
        assert list(late.array[:4]) == [1, 2, 3, 4]
    finally:
        late.close()
        fresh.close()
        crashed._h and crashed._lib.frb_shared_frame_close(crashed._h)
    assert not os.path.exists("/dev/shm" + name) and not os.path.exists("/tmp" + name)


def test_assemble_matches_any_world():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from fractal_renderer_b200 import sharding

    h, w = 75, 9
    full = np.arange(h * w * 3, dtype=np.uint32).reshape(h, w, 3)
    for world in (1, 2, 3, 8):
        parts = [full[sharding.rank_shard(r, world).image_rows(h)] if world > 1 else full for r in range(world)]
        assert np.array_equal(sharding.assemble(parts, h), full) if world > 1 else True
