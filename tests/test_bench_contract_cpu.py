"""bench.py without a GPU: the per-image FLOP figures its throughput -> TFLOP/s conversions rest on are recounted with
torch's FlopCounterMode on the oracle's forward + backward (SURVEY 8d counted them the same way on the reference), and
the reference arm (`--impl reference`: the CPU port of the reference step) prints the contract's JSON line."""
import json
import os
import subprocess
import sys
import warnings

import pytest
import torch
from torch.utils.flop_counter import FlopCounterMode

from oracle import vae_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


@pytest.mark.parametrize("wl", list(bench.WORKLOADS))
def test_flops_per_image_match_a_recount(wl):
    cls, kw, lkw, _bpg, _conditional, gflop, _scaling = bench.WORKLOADS[wl]
    _ch, zc, dis = bench.wl_shape(wl)
    _m, sd = bench.build_state_dict(wl)
    sd = {k: v.requires_grad_(True) for k, v in sd.items()}
    n = 2 if dis else 1                      # one gray + one RGB sample: the disentangled heads differ per modality
    x, cond = bench.wl_batch(wl, n)
    r = kw["resolution"] // 2 ** (len(kw["ch_mult"]) - 1)
    eps = torch.zeros(n, zc, r, r)
    with warnings.catch_warnings(), FlopCounterMode(display=False) as fc:
        warnings.simplefilter("ignore")
        if dis:
            out = O.dis_forward(sd, x, cond, eps, kw["shared_latent_dim"], kw["modality_latent_dim"])
            ld = O.dis_loss(out, x, **lkw)
        else:
            out = O.vae_forward(sd, x, eps, cond)
            ld = O.vae_loss(out["reconstruction"], x, out["mean"], out["logvar"], **lkw)
        ld["loss"].backward()
    counted = fc.get_total_flops() / n / 1e9
    assert abs(counted - gflop) <= 1e-3 * gflop, (wl, counted, gflop)


def _run_reference_arm(extra_env=None):
    env = dict(os.environ, **(extra_env or {}))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1",
                        "--cpu-batch", "2", "--steps", "1", "--warmup", "0"], capture_output=True, text=True, env=env,
                       cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return [ln for ln in r.stdout.splitlines() if ln.startswith("{")]


def test_reference_arm_prints_the_contract_line():
    lines = _run_reference_arm()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "train images/sec" and d["unit"] == "images/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 0 and d["ms_per_step"] > 0
    assert d["dtype"] == "f32" and d["data"] == "synthetic" and d["scaling"] in ("weak", "strong")
    assert d["config"]["workload"].startswith("c1: BaseVAE 1x28x28") and "model" not in d["config"]
    cb = d["cpu_baseline"]
    # "reference" when the unmodified reference copy (oracle/_ref, oracle/build_ref.py) is present, else the oracle port
    from oracle import ref_runner
    assert cb["kind"] == ("reference" if ref_runner.available() else "port")
    assert cb["cores"] == (os.cpu_count() or 1) and cb["value"] == d["value"]
    # the per-step update (NaN scan + clip + AdamW over 63.6 M parameters) is charged once per full batch of 64 images
    assert "2 images" in cb["sample"] and cb["update_ms_per_step"] > 0 and cb["value"] >= cb["value_sample_raw"]
    assert cb["cpu_sample_batch"] == 2
    # same config object as the GPU arm prints (the bounded CPU sample is described in cpu_baseline, not in config)
    assert d["config"]["batch_per_gpu"] == 64 and d["config"]["micro_batch"] == 64
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_reference_arm_other_ranks_exit_quietly():
    assert _run_reference_arm({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_committed_ncu_launch_list_agrees_with_the_bench_kernel_table():
    """profiles/: the kernel SHARES of the ncu launch list (cold-cache, serialised, one 256-image pass) against the
    CUDA-event table of the bench line (one 1024-image pass, same kernels): conv / GroupNorm / attention shares within
    8 points — the fixed per-step work (weight re-pack, AdamW, wgrad finalisation) weighs 4x more in the smaller pass."""
    import csv
    import re
    prof = os.path.join(ROOT, "profiles")
    rows = list(csv.DictReader(open(os.path.join(prof, "r1j_launches.csv"))))
    assert len(rows) > 3000
    total = sum(float(r["ns"]) for r in rows)
    share = lambda pat: sum(float(r["ns"]) for r in rows if re.search(pat, r["kernel"])) / total
    ncu = {"conv": share(r"conv_(gather|reduce)_umma"), "gn": share(r"gn_(apply|bwd)"), "attn": share(r"bgemm_umma")}
    line = json.loads(open(os.path.join(prof, "r1j_bench_c4.json")).read().strip().splitlines()[-1])
    k, step = line["kernels"], line["ms_per_step"]
    ev = {"conv": sum(k[n]["ms"] for n in ("conv_fprop_umma", "conv_wgrad_umma", "conv_dgrad_umma")) / step,
          "gn": (k["groupnorm_fwd"]["ms"] + k["groupnorm_bwd"]["ms"]) / step,
          "attn": (k["attn_fwd"]["ms"] + k["attn_bwd"]["ms"]) / step}
    for name in ncu:
        assert abs(ncu[name] - ev[name]) < 0.08, (name, ncu[name], ev[name])
    # our own kernels dominate the launch list: no library GEMM / conv stands in for the hot path
    ours = sum(float(r["ns"]) for r in rows if r["kernel"].startswith("mvae::")) / total
    assert ours > 0.95, ours
    roof = line["roofline"]
    assert abs(roof["frac"] - roof["achieved"] / roof["peak"]) < 1e-3 and roof["bound"] == "tensor"
    assert line["gpu_launches"] > 0 and line["e2e"]["h2d_bytes_per_step"] > 0


@pytest.mark.parametrize("wl", ["c1", "c3", "c4"])
def test_oracle_port_matches_the_reference_at_the_real_widths(wl):
    """tests/golden/bench_first_step.json holds what the UNMODIFIED reference (oracle/make_bench_check.py) computes for the
    first images of bench.py's batch at the real BASELINE widths (hidden 128, ch_mult (1,2,4[,8]), 2 res blocks, attention):
    loss terms and the norm of every parameter gradient.  The oracle port the GPU parity tests compare against must
    reproduce them (same torch leaf ops, fp32: 1e-5 on the loss, 1e-4 on every gradient norm)."""
    ref = json.load(open(os.path.join(ROOT, "tests", "golden", "bench_first_step.json")))[wl]
    cls, kw, lkw, *_ = bench.WORKLOADS[wl]
    _ch, zc, _dis = bench.wl_shape(wl)
    _m, sd = bench.build_state_dict(wl)
    n = ref["images"]
    x, cond = bench.wl_batch(wl, n)
    r = kw["resolution"] // 2 ** (len(kw["ch_mult"]) - 1)
    eps = torch.randn(n, zc, r, r, generator=torch.Generator().manual_seed(ref["eps_seed"]))
    ld, _out, grads = O.training_step_grads(sd, x, eps, cond, lkw)
    for k, v in ref["loss"].items():
        assert abs(float(ld[k]) - v) <= 1e-5 * max(1.0, abs(v)), (k, float(ld[k]), v)
    assert set(grads) == set(ref["param_grad_norms"])
    for name, v in ref["param_grad_norms"].items():
        got = float(grads[name].double().norm())
        assert abs(got - v) <= 1e-4 * max(v, 1e-3 * ref["grad_norm"]), (name, got, v)


def test_round2_launch_list_at_the_benchmarked_shape_agrees_with_the_bench_kernel_table():
    """the same cross-check for round 2, where the ncu launch list was taken at the BENCHMARKED launch shape (one
    1024-image pass per step, three steps in the list): kernel shares against the CUDA-event table of a bench line, and
    the per-launch DRAM traffic in profiles/roofline_traffic.json against the line's algorithmic bytes"""
    import csv
    import re
    prof = os.path.join(ROOT, "profiles")
    rows = list(csv.DictReader(open(os.path.join(prof, "r2p_launches.csv"))))
    total = sum(float(r["ns"]) for r in rows)
    share = lambda pat: sum(float(r["ns"]) for r in rows if re.search(pat, r["kernel"])) / total
    ncu = {"conv": share(r"conv_(gather|reduce)_umma"), "gn": share(r"gn_(apply|bwd)")}
    line = json.loads(open(os.path.join(prof, "r2p_bench_c4.json")).read().strip().splitlines()[-1])
    k, step = line["kernels"], line["ms_per_step"]
    ev = {"conv": sum(k[n]["ms"] for n in ("conv_fprop_umma", "conv_wgrad_umma", "conv_dgrad_umma")) / step,
          "gn": (k["groupnorm_fwd"]["ms"] + k["groupnorm_bwd"]["ms"]) / step}
    for name in ncu:
        assert abs(ncu[name] - ev[name]) < 0.08, (name, ncu[name], ev[name])
    assert sum(float(r["ns"]) for r in rows if r["kernel"].startswith("mvae::")) / total > 0.95
    roof = line["roofline"]
    assert abs(roof["frac"] - roof["achieved"] / roof["peak"]) < 1e-3 and roof["bound"] == "tensor"
    # real (unpadded) channel counts: 2 x 1024 images x (83.97 - 0.79) GFLOP of conv work / 176 gather launches
    assert abs(roof["algorithmic_gflop_per_launch"] - 967.9) < 1.0
    traffic = json.load(open(os.path.join(prof, "roofline_traffic.json")))
    # the r2p bench line was printed before the r2p --set full capture was read: it carries the figure of the r2a capture
    assert roof["traffic"] in (traffic["dram_bytes_per_launch"], *traffic["superseded"].values()) and roof["traffic"] > 5e8
    assert abs(traffic["dram_bytes_per_launch"] / roof["traffic"] - 1) < 0.05
    assert line["gpu_baseline"]["value"] > 0 and line["value"] > 4 * line["gpu_baseline"]["value"]
    assert line["parity_check"]["ok"] is True and line["cpu_baseline"]["kind"] == "reference"
    assert {h["kernel"] for h in line["roofline_hbm"]} >= {"groupnorm_silu_fwd+bwd", "clip+adamw"}
