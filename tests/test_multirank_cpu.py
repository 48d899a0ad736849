"""world_size-2 tests of the N > 1 host logic on CPU (gloo): KV-head sharding of a workload, the
rank helper bench.py times with (barrier, max over ranks), and that the reference arm runs on
rank 0 only.  The compute stand-in here is the CPU oracle (tests may use it); the product kernels
need a GPU and are covered by the ``-m gpu`` tests and the multi-GPU bench."""
from __future__ import annotations

import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import synth
from tests.harness import oracle_step

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, cfg, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import bench

    R = bench.Rank(type("A", (), {"gpus": world})(), device_type="cpu")
    try:
        w = synth.make_workload(cfg, "cpu")                   # same seed => identical on every rank
        ws = synth.shard_workload(w, rank, world)
        # request tables are rank-independent; pools and q are head slices
        assert torch.equal(ws.req_to_token, w.req_to_token) and torch.equal(ws.out_cache_loc, w.out_cache_loc)
        assert ws.k_pool.shape[1] == cfg.num_kv_heads // world
        out, lse, kp, vp, _, _ = oracle_step(ws)
        out_t = torch.from_numpy(np.ascontiguousarray(out)).float()
        parts = [torch.empty_like(out_t) for _ in range(world)]
        dist.all_gather(parts, out_t)
        kparts = [torch.empty_like(kp) for _ in range(world)]
        dist.all_gather(kparts, kp)
        # max over ranks of a per-rank time: rank r reports r + 1
        mx = R.max_over_ranks(float(rank + 1))
        R.barrier()
        if rank == 0:
            ret["out"] = torch.cat(parts, dim=1).numpy()
            ret["k_pool"] = torch.cat(kparts, dim=1)
            ret["max"] = mx
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
@pytest.mark.parametrize("fp8", [False, True])
def test_kv_head_sharding_equals_global(world, fp8):
    cfg = synth.Config("mr", 5, 0, 16, 4, 64, page_size=4, ragged=(3, 90), share_groups=2, seed=77,
                       kv_dtype="fp8_e4m3" if fp8 else "bf16", k_scale=0.5 if fp8 else 1.0,
                       v_scale=2.0 if fp8 else 1.0)
    w = synth.make_workload(cfg, "cpu")
    ref, _, kp, _, _, _ = oracle_step(w)
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), cfg, ret), nprocs=world, join=True)
        got, k_pool, mx = ret["out"], ret["k_pool"], ret["max"]
    # heads are independent units: the concatenation of the rank slices IS the global result
    assert np.array_equal(got, ref.astype(np.float32))
    assert torch.equal(k_pool, kp)
    assert mx == float(world)


def test_shard_shapes_follow_llama_tp():
    c = synth.CONFIGS["C4"]
    for tp, (hq, hkv) in {1: (64, 8), 2: (32, 4), 4: (16, 2), 8: (8, 1)}.items():
        s = synth.shard(c, tp)
        assert (s.num_q_heads, s.num_kv_heads) == (hq, hkv)
    with pytest.raises(ValueError):
        synth.shard(c, 16)


def test_reference_arm_runs_on_rank0_only():
    """`bench.py --impl reference` under torchrun: ranks != 0 exit 0 without output."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
