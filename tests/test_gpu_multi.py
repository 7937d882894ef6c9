"""Two ranks, one GPU each: the point-sharded solve -- with the small collectives on the
peer-memory all-reduce kernel (csrc/comm.cu) and, separately, with every collective on NCCL -- must reach the cost of the
unsharded solve on one GPU, and the two communication paths must agree with each other to the last bits.  The opt-in
row-sharded PCG (SBA_PCG_SHARDED=1, csrc/pcg_blocks.cu DIST) is checked against the replicated solve to rounding.

Needs two GPUs: skipped on a one-GPU box (the bench lines under profiles/r01_bench_c2_*gpu_weak*.json carry the same
comparison at 2 / 4 / 8 GPUs).  The reference has no distributed path; the oracle for the
sharded solve is the unsharded one.
"""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import importlib, json, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
out = {}
for name, solver in (("C1", m.capi.DENSE_SCHUR), ("C1", m.capi.ITERATIVE_SCHUR)):
    cfg = dict(m.scene.CONFIGS[name])
    prob, _ = m.scene.make_scene(seed=3, **cfg)
    o = m.capi.default_solver_options()
    o.max_num_iterations = 8
    o.linear_solver_type = solver
    obj = [m.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    s = m.Solver(prob, shard_rank=rank, shard_count=world)
    s.init_comm(world, rank, obj[0])
    costs = []
    for _ in range(3):  # the third solve replays CUDA graphs with the collectives inside
        s.reset()
        costs.append(s.solve(o).final_cost)
    single = None
    if rank == 0:
        s1 = m.Solver(prob)
        single = s1.solve(o).final_cost
    out["%s/%d" % (name, solver)] = {"sharded": costs, "single": single}
    dist.barrier()
if rank == 0:
    print("RESULT " + json.dumps(out))
dist.destroy_process_group()
"""


def _run(tmp_path, env_extra):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), ROOT],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-4000:])
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


@pytest.mark.gpu
def test_two_rank_solve_matches_single_gpu_on_both_collective_paths(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    peer = _run(tmp_path, {})                                    # peer-memory all-reduce, replicated PCG (the default)
    dist_pcg = _run(tmp_path, {"SBA_PCG_SHARDED": "1"})          # + the PCG's block rows split across the ranks
    nccl = _run(tmp_path, {"SBA_NO_P2P_ALLREDUCE": "1"})         # everything on NCCL (replicated PCG)
    for key, res in peer.items():
        single = res["single"]
        for run in (res, dist_pcg[key], nccl[key]):
            for c in run["sharded"]:
                assert abs(c - single) <= 1e-7 * single, (key, c, single)   # north-star bar is 1e-6
            assert max(run["sharded"]) - min(run["sharded"]) <= 1e-12 * single, (key, run)  # plain launches == graph replay
        # two ranks: a sum of two terms is the same in either order, so the two collective paths agree bit for bit
        assert res["sharded"][-1] == nccl[key]["sharded"][-1], (key, res, nccl[key])
        # the row-sharded PCG adds its scalars rank by rank instead of CTA by CTA: same solve to rounding
        assert abs(dist_pcg[key]["sharded"][-1] - res["sharded"][-1]) <= 1e-9 * single, (key, dist_pcg[key], res)
