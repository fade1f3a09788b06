"""Host-side logic of the N > 1 benchmark path on CPU: two gloo ranks reproduce bench.py's
aggregation (barrier, max over ranks of the per-rank time, sum of the per-rank work) and the
`--impl reference` rule that only rank 0 prints.  The measured path itself shards with no collective
(one independent engine per GPU), so this is all the cross-rank logic there is."""
import json
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_aggregation_with_gloo(tmp_path):
    script = tmp_path / "ranks.py"
    script.write_text(textwrap.dedent("""
        import json, os, sys
        import torch, torch.distributed as dist
        dist.init_process_group("gloo")
        rank, world = dist.get_rank(), dist.get_world_size()
        dist.barrier()
        dev_ms = torch.tensor([100.0 + 50.0 * rank], dtype=torch.float64)     # per-rank device time
        launches = torch.tensor([81.0 * 255], dtype=torch.float64)
        dist.all_reduce(dev_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(launches, op=dist.ReduceOp.SUM)
        tokens_all = 256 * 1 * world
        if rank == 0:
            print(json.dumps({"value": tokens_all / (dev_ms.item() / 1e3), "n_gpus": world,
                              "gpu_launches": int(launches.item())}))
        dist.destroy_process_group()
    """))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                                     # rank 0 alone prints
    d = json.loads(lines[0])
    assert d["n_gpus"] == 2 and d["gpu_launches"] == 2 * 81 * 255
    assert abs(d["value"] - 512 / 0.150) < 1e-6                # total tokens / max-over-ranks time


def test_reference_arm_only_rank0_prints():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
