"""CPU: the node reordering of the plug-in's graph_optimize hook (backend/ggml-b200.cpp) on cgraphs built with the reference's ggml in
llama.cpp's node order — still a valid order, q / k / v and gate / up adjacent, nothing moved across an in-place op on the shared
activations (oracle/reorder_check.cpp; needs oracle/_ref and the built plug-in, no device)."""
import subprocess
from pathlib import Path

import pytest

from oracle import refenv

ROOT = Path(__file__).resolve().parents[1]


def test_graph_optimize_reordering_is_valid_and_groups_shared_activations():
    exe = refenv.REF_DIR / "reorder_check"
    plugin = ROOT / "llama.cpp-s390x_b200" / "libggml-b200.so"
    if not exe.exists() or not plugin.exists():
        pytest.skip("oracle/_ref/reorder_check or the plug-in is not built")
    p = subprocess.run([str(exe), str(plugin)], capture_output=True, text=True, timeout=120, env=refenv.ref_env())
    assert p.returncode == 0 and p.stdout.strip().endswith("OK"), p.stdout + p.stderr
    assert "6 groups adjacent" in p.stdout and "in-place pin: order kept" in p.stdout
