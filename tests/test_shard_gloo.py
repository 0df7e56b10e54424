"""Multi-GPU path on CPU: world_size-2 gloo job that shards streams as bench.py does and checks that the
shards tile the stream set with no overlap and that the max-over-ranks reduction works."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_properties():
    from audio_codec_b200.shard import shard_range
    for n in (1, 7, 4096, 100000):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_two_rank_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys, torch, torch.distributed as dist
        sys.path.insert(0, {ROOT!r})
        from audio_codec_b200.shard import shard_range
        dist.init_process_group("gloo")
        r, w = dist.get_rank(), dist.get_world_size()
        lo, hi = shard_range(4099, r, w)
        cnt = torch.tensor([hi - lo], dtype=torch.int64)
        dist.all_reduce(cnt)
        t = torch.tensor([1.0 + r], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert cnt.item() == 4099 and t.item() == float(w), (cnt, t)
        dist.barrier()
        dist.destroy_process_group()
    """))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
