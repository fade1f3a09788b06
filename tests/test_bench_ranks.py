"""Host-side logic of the N > 1 benchmark path on CPU: two gloo ranks run bench.py's own
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
    """bench.py's own Ranks / aggregate code (imported, not restated) on two gloo ranks with made-up per-rank timings."""
    script = tmp_path / "ranks.py"
    script.write_text(textwrap.dedent("""
        import json, os, sys
        sys.path.insert(0, %r)
        import bench
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        ranks = bench.Ranks(world, backend="gloo")
        ranks.barrier()
        # rank r: prefill 10 + 5 r ms, decode 90 + 45 r ms, wall 0.2 + 0.1 r s, 81 * 255 launches, 1 step
        agg = bench.aggregate(ranks, 10.0 + 5.0 * rank, 90.0 + 45.0 * rank, 0.2 + 0.1 * rank, 81 * 255, 1)
        # the host-side hand-shake that keeps the other ranks off their GPUs while rank 0 runs its extra legs
        import time
        t0 = time.time()
        if rank == 0:
            time.sleep(0.5)
            ranks.host_signal("extra_legs_done")
        else:
            ranks.host_wait("extra_legs_done")
            assert time.time() - t0 >= 0.4
        if rank == 0:
            print(json.dumps(dict(agg, n_gpus=world)))
        ranks.close()
    """ % ROOT))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                                     # rank 0 alone prints
    d = json.loads(lines[0])
    assert d["n_gpus"] == 2 and d["launches_all"] == 2 * 81 * 255
    assert d["dev_ms"] == 150.0 and d["pre_ms"] == 15.0 and d["dec_ms"] == 135.0      # max over ranks
    assert abs(d["value"] - 512 / 0.150) < 1e-6                # total tokens / max-over-ranks device time
    assert abs(d["e2e"] - 512 / 0.3) < 1e-6                    # total tokens / max-over-ranks wall time


def test_reference_arm_only_rank0_prints():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
