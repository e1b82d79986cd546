"""The two rules the fused tree kernel of the builder (rtb_build.cu: k_fit_collapse / k_fit_top_fused) rests on, checked on the CPU
against a top-down construction of the same radix tree (tools/apetrei_model.py): the bottom-up parent rule, and the test that
tells which nodes lie inside a chunk of consecutive leaves."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bottom_up_radix_tree_and_chunk_locality_rules():
    for seed in ("0", "7"):
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "apetrei_model.py"), seed], capture_output=True, text=True, timeout=300)
        assert out.returncode == 0 and out.stdout.strip() == "ok", out.stdout[-500:] + out.stderr[-2000:]
