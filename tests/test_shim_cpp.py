"""The C++ shim (shim/sba_bundle_adjuster.h) replays the reference's BundleAdjuster test scenarios
(src/optim/bundle_adjustment_test.cc:186-645) over the mock Reconstruction: host part on CPU, solve part on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "semantic-bundle-adjustment-colmap_b200")
EXE = os.path.join(ROOT, "tests", "shim_test.bin")


@pytest.fixture(scope="module")
def shim_exe(sba):
    sba.capi.load()  # the C-ABI library must exist (build() made it); never built lazily here
    src = os.path.join(ROOT, "tests", "shim_test.cpp")
    hdrs = [os.path.join(PKG, "shim", h) for h in ("sba_bundle_adjuster.h", "sba_semantic_adjusters.h", "sba_mapper_passes.h", "sba_rig_bundle_adjuster.h", "mock_reconstruction.h")]
    lib = os.path.join(PKG, "libsba_b200.so")
    newest = max(os.path.getmtime(p) for p in [src, lib] + hdrs)
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < newest:
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror",
                               "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(PKG, "shim"), src, "-o", EXE,
                               "-L" + PKG, "-lsba_b200", "-Wl,-rpath,$ORIGIN/../semantic-bundle-adjustment-colmap_b200"])
    return EXE


def test_shim_counts_and_errors(shim_exe):
    out = subprocess.run([shim_exe, "counts"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr


@pytest.mark.gpu
def test_shim_solve_on_device(shim_exe):
    out = subprocess.run([shim_exe, "solve"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "two_view: cost" in out.stdout and "semantic: cost" in out.stdout and "cylinder: cost" in out.stdout
    assert "refine_absolute_pose: cost" in out.stdout and out.stdout.count("rig ") == 4


def test_flatten_bench_helper_reports_json(sba):
    """tools/shim_flatten_bench.bin (built by __graft_entry__.build()) -- the helper behind bench.py's `host_flatten` key."""
    import json
    exe = os.path.join(ROOT, "tools", "shim_flatten_bench.bin")
    if not os.path.exists(exe):
        pytest.skip("helper not built (run __graft_entry__.build())")
    out = subprocess.run([exe, "20", "3000", "--json"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["images"] == 20 and d["points"] == 3000 and d["observations"] > 2 * 3000 and d["flatten_ms"] > 0
