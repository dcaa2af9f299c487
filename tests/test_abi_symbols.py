"""CPU: the C-ABI library loads and exports every symbol include/ldso_gpu.h declares (no compute calls),
and the oracle mirrors each of them under the orc_ prefix."""
import ctypes
import os
import re

from ldso_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ldso_gpu.h")).read() + open(os.path.join(ROOT, "include", "ldso_system.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ldso_gpu_[a-z0-9_]+)\s*\(", src)))


def test_gpu_library_exports_every_declared_symbol():
    build.build_gpu()
    lib = ctypes.CDLL(build.GPU_LIB)
    syms = declared_symbols()
    assert len(syms) >= 15
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_oracle_mirrors_the_abi():
    build.build_oracle()
    lib = ctypes.CDLL(build.ORACLE_LIB)
    skip = {"ldso_gpu_timer_start", "ldso_gpu_timer_stop", "ldso_gpu_profile_enable", "ldso_gpu_profile_get",
            "ldso_gpu_tracker_calc_res_noalloc"}     # device-memory plumbing with no meaning on the CPU
    missing = [s for s in declared_symbols() if s not in skip and not hasattr(lib, s.replace("ldso_gpu_", "orc_"))]
    assert not missing, missing


def test_config_struct_layout_matches_header():
    api = capi.Api(build.ORACLE_LIB, "orc_")
    cfg = api.default_config()
    assert (cfg.w, cfg.h) == (640, 480) and cfg.huber_th == 9 and cfg.coarse_cutoff_th == 20
    assert cfg.frame_energy_th_const_weight == 0.5 and cfg.tracker_solve_mode == 0   # last fields: no layout drift
    gapi = capi.Api(build.GPU_LIB, "ldso_gpu_")
    g = gapi.default_config()
    assert bytes(g) == bytes(cfg)


def test_product_never_references_the_oracle():
    pkg = os.path.join(ROOT, "ldso_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp")) and f != "build.py":
                txt = open(os.path.join(dp, f)).read()
                assert "oracle/" not in txt and "liboracle" not in txt and "orc_" not in txt, f
