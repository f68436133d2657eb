"""bench.py without a GPU: the byte model behind `roofline` equals the table of DESIGN.md section 5, the command line keeps
the driver's contract (defaults N = 1, warm-up >= 3, flags of both arms), and the committed records carry every key the
contract names."""
import glob
import json
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_byte_model_matches_design_table():
    import bench
    N, E, H, L = 98304, 3_900_000, 128, 2
    assert bench.algorithmic_bytes("tn_msg_fwd", N, E, H, L) == 72 * H * N + 8 * E
    assert bench.algorithmic_bytes("tn_msg_bwd", N, E, H, L) == 144 * H * N + 16 * E
    assert bench.algorithmic_bytes("tn_embed_edge_fwd", N, E, H, L) == 36 * H * N + 20 * E
    assert bench.algorithmic_bytes("tn_embed_edge_bwd", N, E, H, L) == 36 * H * N + 20 * E + 12 * N
    assert bench.algorithmic_bytes("neighbors_csr", N, E, H, L) == 16 * N + 20 * E
    assert bench.algorithmic_bytes("et_edge_fwd", N, E, H, L) == 48 * H * N + 20 * E
    assert bench.algorithmic_bytes("et_edge_bwd", N, E, H, L) == 96 * H * N + 40 * E
    assert bench.algorithmic_bytes("node_gemm_tc", N, E, H, L) == 72 * H * N          # everything else: read + write a node tensor
    assert [bench.l2_bytes(f, E, H) // (H * E) for f in ("tn_msg_fwd", "tn_msg_bwd", "tn_embed_edge_fwd", "et_edge_fwd",
                                                         "et_edge_bwd")] == [84, 120, 48, 92, 112]
    design = open(os.path.join(ROOT, "DESIGN.md")).read()
    for cell in ("`72·H·N + 8·E`", "`144·H·N + 16·E`", "`84·H·E`", "`120·H·E`", "`16·N + 20·E`", "`92·H·E` / `112·H·E`"):
        assert cell in design, cell


def test_command_line_contract():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0
    text = out.stdout + out.stderr          # bench.py keeps stdout for the one JSON line; everything else goes to stderr
    for flag in ("--gpus", "--steps", "--warmup", "--impl", "--workload", "--no-strong", "--transport"):
        assert flag in text, flag
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert re.search(r'"--gpus", type=int, default=1\b', src)
    assert re.search(r'"--impl", default="b200", choices=\["b200", "reference"\]', src)
    m = re.search(r'"--warmup", type=int, default=(\d+)', src)
    assert m and int(m.group(1)) >= 3


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "profiles", "r03*_bench_*.json"))))
def test_committed_records_carry_the_contract_keys(path):
    d = json.load(open(path))
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "e2e", "gpu_launches", "cpu_baseline"):
        assert k in d, (path, k)
    assert d["unit"] == "atom-steps/s" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"])
    if d.get("impl") == "reference":
        assert d["gpu_launches"] == 0 and d["cpu_baseline"]["kind"] in ("port", "reference")
        assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
        return
    assert d["gpu_launches"] > 0 and "workload" in d["config"] and "model" not in d["config"]
    r = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r) and r["bound"] in ("hbm", "tensor")
    assert abs(r["frac"] - r["achieved"] / r["peak"]) <= 1e-9
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    assert not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(d["clocks"]["reasons"])
    if d.get("strong_scaling"):
        ss = d["strong_scaling"]
        assert ss["atoms_total"] == 1_029_000 and ss["n_gpus"] == d["n_gpus"]
        assert ss["correctness"]["max_abs_dF"] <= 1e-4 and ss["correctness"]["rel_dE"] <= 1e-5


def test_reference_arm_under_torchrun_on_the_cpu():
    """`bench.py --impl reference` launched the way the driver launches N > 1 (torchrun, one process per GPU): rank 0 alone
    times the CPU restatement and prints ONE JSON line, the other rank exits 0 without work; no GPU is touched."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29577", os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
           "--warmup", "1"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["steps"] == 1 and d["warmup"] == 1
    assert d["unit"] == "atom-steps/s" and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert abs(d["ms_per_step"] * 1e-3 * d["value"] - 166) <= 1e-6 * 166          # value = atoms / seconds per step
