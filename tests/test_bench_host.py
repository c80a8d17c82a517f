"""CPU: the host-side bookkeeping behind bench.py's roofline keys (no GPU, no timing).
* the executed / algorithmic flop counts of the FD kernels are the figures DESIGN.md section 4.1 and SURVEY.md 8(d) state;
* profiles/traffic.json speaks for the CURRENT kernel sources (bench.py reports roofline.traffic = null otherwise, so a kernel
  edit without a fresh ncu capture shows up here before it shows up as a null in the driver's line);
* profiles/j2_flops.json carries the keys the config-5 roofline evaluates;
* the slab cut used for strong scaling covers every layer exactly once for any rank count."""
import json
import os

import numpy as np
import pytest

import bench
from mimi_b200 import partition as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_flop_model_matches_the_stated_figures():
    alg, exe, fd = bench.flops_per_qp(2, 3, "neohookean")
    nq, tdof = 64, 81
    # 1 189 flop per qp per evaluation, 82 evaluations; VERDICT's hand check: 1 188 + 81 * 862 + 2 * 81^2 / 64
    assert alg == 82 * 1189
    assert exe == 1188 + 81 * 862
    assert fd == 2 * tdof * tdof
    assert abs((exe + fd / nq) - 71215.03) < 0.01
    alg3, exe3, fd3 = bench.flops_per_qp(3, 3, "stvk")
    assert alg3 > exe3 > 0 and fd3 == 2 * 192 * 192


@pytest.mark.parametrize("key", ["element_fd_kernel<3,27,64,neohookean>", "element_fd_split_kernel<3,64,125,stvk,4>",
                                 "element_fast_p2v3_kernel<neohookean>"])
def test_committed_ncu_traffic_is_of_the_current_kernel_source(key):
    T = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    assert key in T, "no ncu capture recorded for %s (scripts/ncu_traffic.py)" % key
    assert T[key]["source_md5"] == bench.source_md5(key), (
        "%s was edited after its ncu --set full capture: re-capture it and run scripts/ncu_traffic.py" % key)
    traffic, note = bench.ncu_traffic(key, T[key]["launch_elements"])
    assert traffic == T[key]["dram_read_bytes"] + T[key]["dram_write_bytes"]


def test_j2_flop_model_file():
    J = json.load(open(os.path.join(ROOT, "profiles", "j2_flops.json")))
    for k in ("flop_fixed_per_qp", "flop_per_hardening_eval", "fused_share_fixed", "fused_share_variable", "capture"):
        assert k in J
    assert 5e4 < J["flop_fixed_per_qp"] < 1e5 and 100 < J["flop_per_hardening_eval"] < 1000
    # the model reproduces the capture it was calibrated on
    assert abs(J["flop_fixed_per_qp"] + J["flop_per_hardening_eval"] * J["hardening_evals_per_qp"] - J["flop_per_qp"]) <= 1e-6 * J["flop_per_qp"]


@pytest.mark.parametrize("layers,world", [(100, 8), (100, 3), (7, 7), (64, 4), (13, 2)])
def test_slab_bounds_cover_every_layer_once(layers, world):
    kb = P.slab_bounds(layers, world)
    assert kb[0] == 0 and kb[-1] == layers and len(kb) == world + 1
    d = np.diff(kb)
    assert d.min() >= layers // world and d.max() <= layers // world + 1 and (np.diff(d) <= 0).all()
