"""GPU conformance (-m gpu): the reference's own UNMODIFIED harness, tests/test-backend-ops.cpp (compiled from
/root/reference into oracle/_ref by oracle/Makefile), run against our backend plug-in loaded through GGML_BACKEND_PATH
(ggml/src/ggml-backend-reg.cpp:581-585).  It compares every MUL_MAT / MUL_MAT_ID case our supports_op() accepts with the
reference CPU backend (NMSE <= 5e-4) and checks the sentinels around every tensor for out-of-bounds writes."""
import re
import subprocess
from pathlib import Path

import pytest

from oracle import refenv

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
PLUGIN = ROOT / "llama.cpp-s390x_b200" / "libggml-b200.so"
HARNESS = refenv.REF_DIR / "test-backend-ops"


def run_harness(*args, timeout=900, extra_env=None):
    if not HARNESS.exists() or not PLUGIN.exists():
        pytest.skip("oracle/_ref/test-backend-ops or libggml-b200.so not built (needs /root/reference at build time)")
    env = refenv.ref_env({"GGML_BACKEND_PATH": str(PLUGIN), **(extra_env or {})})
    p = subprocess.run([str(HARNESS), *args], env=env, capture_output=True, text=True, timeout=timeout)
    out = re.sub(r"\x1b\[[0-9;]*m", "", p.stdout + p.stderr)
    return p.returncode, out


@pytest.mark.parametrize("op", ["MUL_MAT", "MUL_MAT_ID"])
def test_backend_ops(op):
    rc, out = run_harness("test", "-b", "B2000", "-o", op)
    tail = "\n".join(out.splitlines()[-25:])
    assert "Backend B2000" in out, tail                     # the plug-in was loaded and selected
    m = re.search(r"(\d+)/(\d+) tests passed", out)
    assert m, tail
    passed, total = int(m.group(1)), int(m.group(2))
    assert rc == 0 and passed == total and total > 0, tail
    # every case of the four weight types with F32 activations must have RUN (not "not supported")
    ran = [l for l in out.splitlines() if re.search(r"type_a=(q4_0|q8_0|q4_K|q6_K),type_b=f32", l) and l.rstrip().endswith("OK")]
    assert len(ran) >= (20 if op == "MUL_MAT" else 20), f"only {len(ran)} quantized cases ran\n{tail}"
    failed = [l for l in out.splitlines() if "FAIL" in l]
    assert not failed, failed[:5]


def test_backend_ops_more_weight_types():
    """Q4_1 / Q5_0 / Q5_1 / Q5_K / Q2_K / Q3_K / IQ4_NL / IQ4_XS / MXFP4 / NVFP4 weights through the unmodified harness — every MUL_MAT case of
    those types with F32 activations must run and match the CPU backend"""
    rc, out = run_harness("test", "-b", "B2000", "-o", "MUL_MAT")
    assert rc == 0 and "Backend B2000: OK" in out, "\n".join(out.splitlines()[-25:])
    for t, least in (("q4_1", 40), ("q5_0", 10), ("q5_1", 10), ("q5_K", 10), ("q2_K", 10), ("q3_K", 10), ("iq4_nl", 10), ("iq4_xs", 10), ("mxfp4", 40), ("nvfp4", 40)):
        ran = [l for l in out.splitlines() if re.search(r"type_a=" + t + r",type_b=f32", l) and l.rstrip().endswith("OK")]
        assert len(ran) >= least, f"only {len(ran)} {t} cases ran"


def test_backend_ops_support_listing():
    rc, out = run_harness("support", "-b", "B2000", "-o", "MUL_MAT")
    assert rc == 0 and "B2000" in out


GRAPH_CHECK = refenv.REF_DIR / "graph_check"


# the other ops of a Llama graph and the fused MUL_MAT patterns (N2): op filter -> least number of cases that must RUN
EXTRA_OPS = {
    "ADD": 40, "SUB": 40, "MUL": 40, "DIV": 40, "ADD_ID": 30, "SCALE": 4, "RMS_NORM": 20, "ROPE": 150, "SOFT_MAX": 150, "CPY": 50, "CONT": 20, "DUP": 6,
    "GET_ROWS": 30, "SET_ROWS": 50, "SILU": 2, "GELU": 2, "RELU": 2, "SIGMOID": 2, "TANH": 2, "SWIGLU": 12, "GEGLU": 12, "REGLU": 12, "GEGLU_ERF": 12,
    "GEGLU_QUICK": 12, "SWIGLU_OAI": 8,
    # whole-graph cases of the harness (run_whole_graph): the plug-in computes every node, fused where its shell fuses
    "MUL_MAT_VEC_FUSION": 100, "MUL_MAT_ID_FUSION": 6, "RMS_NORM_MUL_ADD": 30, "ADD_RMS_NORM": 25, "RMS_NORM_MUL_ROPE": 70, "ROPE_SET_ROWS": 24,
}


@pytest.mark.parametrize("op", sorted(EXTRA_OPS))
def test_backend_ops_llama_graph_ops(op):
    """unmodified test-backend-ops, op by op: everything supports_op() accepts must match the reference's CPU backend (the harness' own
    per-op NMSE limits), and at least the listed number of cases must have run rather than been declined"""
    rc, out = run_harness("test", "-b", "B2000", "-o", op, extra_env={"GGML_B200_FUSE": "2"})       # (2: fuse whatever the pattern allows)
    tail = "\n".join(out.splitlines()[-25:])
    assert rc == 0 and "Backend B2000: OK" in out, tail
    ran = [l for l in out.splitlines() if re.match(r"\s*" + op + r"\(", l) and l.rstrip().endswith("OK")]
    assert len(ran) >= EXTRA_OPS[op], f"only {len(ran)} cases of {op} ran\n{tail}"
    assert not [l for l in out.splitlines() if "FAIL" in l]


@pytest.mark.parametrize("mode", ["on_device", "mul_mat_only"])
@pytest.mark.parametrize("wtype,n_tokens", [("q4_K", 1), ("q4_K", 5), ("q4_K", 64), ("q8_0", 1), ("q8_0", 48), ("q4_0", 1), ("q4_0", 33), ("q6_K", 1), ("q6_K", 20)])
def test_layer_graph_through_sched(wtype, n_tokens, mode):
    """oracle/graph_check.cpp: transformer-layer-shaped graphs scheduled by the reference's ggml_backend_sched over {plug-in, CPU} == the CPU
    backend alone.  on_device: the plug-in takes every node (one split); at one token gate / up / SWIGLU go through b200_mul_mat_fused, wq / wk /
    wv through one b200_mul_mat_batch.  mul_mat_only (GGML_B200_OPS=mul_mat, the round-1 behaviour): all MUL_MATs on the plug-in, everything
    else on the CPU, the scheduler copying activations across; wq/wk/wv and gate/up each one batch."""
    if not GRAPH_CHECK.exists() or not PLUGIN.exists():
        pytest.skip("oracle/_ref/graph_check or libggml-b200.so not built (needs /root/reference at build time)")
    extra = {"GGML_BACKEND_PATH": str(PLUGIN), "GGML_B200_STATS": "1"}
    if mode == "mul_mat_only":
        extra["GGML_B200_OPS"] = "mul_mat"
    p = subprocess.run([str(GRAPH_CHECK), wtype, str(n_tokens), "3"], env=refenv.ref_env(extra), capture_output=True, text=True, timeout=300)
    out = p.stdout + p.stderr
    assert p.returncode == 0 and "OK" in p.stdout.splitlines()[-1], out[-2000:]
    m = re.search(r"stats: graphs=(\d+) mul_mat=(\d+) mul_mat_id=(\d+) fused_groups=(\d+) other_ops=(\d+) fused_mul_mat=(\d+)", out)
    assert m, out[-2000:]
    mul_mat, groups, other, fused = int(m.group(2)), int(m.group(4)), int(m.group(5)), int(m.group(6))
    if mode == "mul_mat_only" or n_tokens > 1:
        # 3 layers x 7 matmuls x 2 passes; two batched groups per layer and pass
        assert mul_mat == 3 * 7 * 2 and groups == 3 * 2 * 2 and fused == 0, m.group(0)
    else:
        # one token: gate + up + SWIGLU is one fused launch per layer and pass, the other five matmuls go through the batch / single path
        assert mul_mat == 3 * 5 * 2 and fused == 3 * 2 and groups == 3 * 2, m.group(0)
    assert (other == 0) == (mode == "mul_mat_only"), m.group(0)


COMM_CHECK = refenv.REF_DIR / "comm_check"


@pytest.mark.parametrize("args", [("2", "4096"), ("2", "1000003"), ("3", "65536", "1"), ("4", "262147"), ("8", "4096", "5"), ("1", "777")])
def test_comm_allreduce_through_the_registry(args):
    """oracle/comm_check.cpp: ggml_backend_comm_init / _allreduce_tensor / _free obtained from the plug-in's registry the way the reference's
    meta (tensor-parallel) backend does: every participant ends with the bit-exact participant-order sum, a tensor
    that was not computed contributes zeros, an F16 tensor is declined.  With one GPU the participants are backends of the same device."""
    if not COMM_CHECK.exists() or not PLUGIN.exists():
        pytest.skip("oracle/_ref/comm_check or libggml-b200.so not built (needs /root/reference at build time)")
    p = subprocess.run([str(COMM_CHECK), *args], env=refenv.ref_env({"GGML_BACKEND_PATH": str(PLUGIN)}), capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and p.stdout.strip().splitlines()[-1] == "OK", (p.stdout + p.stderr)[-2000:]


LLAMA_STEP = refenv.REF_DIR / "llama_step"


@pytest.mark.parametrize("args", [("tiny", "1", "40"), ("tiny", "7", "33"), ("tiny", "64", "0"), ("1b", "1", "300"), ("1b", "33", "100")])
def test_whole_llama_graph_on_the_plugin(args):
    """oracle/llama_step.cpp: a Llama decoder graph as llama.cpp builds it without flash attention (GET_ROWS, RMS_NORM, MUL, quantized MUL_MAT,
    ROPE, SET_ROWS / CPY into an F16 KV cache, F16 MUL_MAT over strided cache views with grouped-query broadcast, SOFT_MAX with mask, CONT, ADD,
    SWIGLU) scheduled by the reference's ggml_backend_sched over {plug-in, CPU}: every node must run on the plug-in (ONE split) and the logits must
    match the reference's CPU backend"""
    if not LLAMA_STEP.exists() or not PLUGIN.exists():
        pytest.skip("oracle/_ref/llama_step or libggml-b200.so not built (needs /root/reference at build time)")
    p = subprocess.run([str(LLAMA_STEP), *args, "2", "1"], env=refenv.ref_env({"GGML_BACKEND_PATH": str(PLUGIN)}), capture_output=True, text=True, timeout=600)
    out = p.stdout + p.stderr
    assert p.returncode == 0 and p.stdout.strip().splitlines()[-1] == "OK", out[-2000:]
    assert '"splits_b200": 1,' in out, out[-2000:]


SPLIT_CHECK = refenv.REF_DIR / "split_check"


@pytest.mark.parametrize("args,shards", [(("q4_K", "4096", "4096", "1", "0.5"), 2), (("q4_K", "1536", "2048", "4", "0.3"), 3), (("q8_0", "1000", "1024", "5", "0.6"), 2),
                                         (("q6_K", "1280", "2048", "1", "0.5"), 4), (("q4_0", "2048", "512", "8", "0.5"), 2), (("q4_K", "2048", "4096", "64", "0.4"), 2),
                                         (("q8_0", "4096", "4096", "512", "0.5"), 2), (("q4_K", "2048", "8192", "512", "0.5"), 4)])
def test_split_buffer_type(args, shards):
    """oracle/split_check.cpp: "ggml_backend_split_buffer_type" from the registry (llama.cpp -sm row), weights scattered by
    set_tensor, MUL_MAT with src1 / dst on the main device == the unsplit result bit for bit and == the CPU backend within
    NMSE 5e-4; the llama.cpp weight placement probe passes; MUL_MAT_ID with split weights is rejected.  The shards are emulated
    on the one GPU of the test box (GGML_B200_SPLIT_EMULATE); on a multi-GPU box the same binary runs over real devices."""
    if not SPLIT_CHECK.exists() or not PLUGIN.exists():
        pytest.skip("oracle/_ref/split_check or libggml-b200.so not built (needs /root/reference at build time)")
    env = refenv.ref_env({"GGML_BACKEND_PATH": str(PLUGIN), "GGML_B200_SPLIT_EMULATE": str(shards)})
    p = subprocess.run([str(SPLIT_CHECK), *args], env=env, capture_output=True, text=True, timeout=300)
    out = p.stdout + p.stderr
    assert p.returncode == 0 and "OK" in p.stdout.splitlines()[-1], out[-2000:]
    assert "buft=B200_Split" in out and "split==plain:1" in out, out[-2000:]


E2E_PLUGIN = refenv.REF_DIR / "e2e_plugin"


def test_e2e_through_the_plugin_small():
    """oracle/e2e_plugin.cpp (bench.py's e2e leg) on a tiny model: set_async -> graph_compute_async -> get_async -> synchronize
    through libggml-b200.so, pinned host memory from the plug-in's host buffer type, spot check against the CPU backend"""
    if not E2E_PLUGIN.exists() or not PLUGIN.exists():
        pytest.skip("oracle/_ref/e2e_plugin or libggml-b200.so not built (needs /root/reference at build time)")
    import json
    env = refenv.ref_env({"GGML_BACKEND_PATH": str(PLUGIN)})
    p = subprocess.run([str(E2E_PLUGIN), "tiny", "3", "1"], env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["n_matmuls"] == 15 and d["nmse_vs_cpu_backend"] <= 5e-4
    assert d["caps"] == {"async": 1, "host_buffer": 1, "events": 1} and d["host_memory"] == "B200_Host"
