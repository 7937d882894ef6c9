"""The C-ABI library: loads, exports every symbol the header declares, validates arguments,
fails loudly without a device; sharding by point ownership; world_size-2 gloo reduction identity."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from util import oracle_evaluate

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    h = open(os.path.join(ROOT, "include", "sba_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(sba_[a-z_0-9]+)\s*\(", h)))


def test_header_symbols_are_exported_and_typed(sba):
    lib = sba.capi.load()
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), "libsba_b200.so does not export " + n
        assert n in sba.capi.SYMBOLS, "ctypes binding lacks " + n
    assert set(sba.capi.SYMBOLS) == set(names)
    assert b"sm_100a" in lib.sba_version()


def test_struct_sizes_match_header(sba):
    # sizeof computed by the C compiler for the header's structs == ctypes mirror
    src = '#include "sba_b200.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu",sizeof(sba_problem_desc),' \
          'sizeof(sba_solver_options),sizeof(sba_iteration_summary),sizeof(sba_summary),sizeof(sba_semantic_desc),sizeof(sba_cylinder_desc),' \
          'sizeof(sba_semantic_problem),sizeof(sba_cylinder_problem),sizeof(sba_rig_desc),sizeof(sba_rig_problem));}'
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "s.c"), "-o", os.path.join(d, "s")])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(d, "s")]).split()]
    c = sba.capi
    assert sizes == [C.sizeof(c.ProblemDesc), C.sizeof(c.SolverOptions), C.sizeof(c.IterationSummary),
                     C.sizeof(c.Summary), C.sizeof(c.SemanticDesc), C.sizeof(c.CylinderDesc), C.sizeof(c.SemanticProblem),
                     C.sizeof(c.CylinderProblem), C.sizeof(c.RigDesc), C.sizeof(c.RigProblem)]


def test_default_options_are_the_reference_defaults(sba):
    o = sba.capi.default_solver_options()  # bundle_adjustment.h:80-93 + Ceres 2.1.0 defaults
    assert (o.max_num_iterations, o.max_linear_solver_iterations, o.max_num_consecutive_invalid_steps) == (100, 200, 10)
    assert (o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance) == (0.0, 0.0, 0.0)
    assert (o.initial_trust_region_radius, o.min_relative_decrease, o.eta, o.jacobi_scaling) == (1e4, 1e-3, 0.1, 1)


def test_no_device_is_an_error_not_a_fallback(sba):
    lib = sba.capi.load()
    if lib.sba_device_count() > 0:
        pytest.skip("a CUDA device is present")
    prob, _ = sba.scene.make_scene(n_cam=3, n_pt=20, mean_track=3.0)
    with pytest.raises(sba.capi.SbaError) as e:
        sba.Solver(prob)
    assert e.value.code == sba.capi.ERR_NO_DEVICE and "no CPU fallback" in str(e.value)


def test_argument_validation(sba):
    prob, _ = sba.scene.make_scene(n_cam=3, n_pt=20, mean_track=3.0)
    bad = sba.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, prob.camera_model, prob.camera_params,
                          prob.points, prob.obs_image, prob.obs_point + 1000, prob.obs_xy)
    with pytest.raises(sba.capi.SbaError) as e:
        bad.counts()
    assert e.value.code == sba.capi.ERR_INVALID_ARGUMENT
    bad = sba.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, prob.camera_model * 0 + 77,
                          prob.camera_params, prob.points, prob.obs_image, prob.obs_point, prob.obs_xy)
    with pytest.raises(sba.capi.SbaError) as e:
        bad.counts()
    assert e.value.code == sba.capi.ERR_UNKNOWN_CAMERA_MODEL


@pytest.mark.parametrize("nranks", [1, 2, 3, 8])
def test_shard_points_partitions_and_balances(sba, nranks):
    prob, _ = sba.scene.make_scene(n_cam=12, n_pt=3000, mean_track=6.0, seed=5)
    pk = np.zeros((nranks, len(prob.points)), np.uint8)
    ok = np.zeros((nranks, prob.num_observations), np.uint8)
    for r in range(nranks):
        pk[r], ok[r] = prob.shard_masks(nranks, r)
    assert (pk.sum(0) == 1).all() and (ok.sum(0) == 1).all()          # a partition
    for r in range(nranks):
        assert (ok[r] == pk[r][prob.obs_point]).all()                  # observations follow their point
        idx = np.nonzero(pk[r])[0]
        assert len(idx) == 0 or (np.diff(idx) == 1).all()              # contiguous point ranges
    per = ok.sum(1)
    assert per.max() - per.min() <= 64                                 # balanced by observation count


def test_empty_and_ragged_inputs(sba):
    e = sba.FlatProblem(np.zeros((0, 4)), np.zeros((0, 3)), [], [], np.zeros((0, 12)), np.zeros((0, 3)), [], [], np.zeros((0, 2)))
    s = e.counts()
    assert s.num_residuals == 0 and s.num_effective_parameters_reduced == 0
    # points without observations and images without observations are not part of the reduced program
    prob, _ = sba.scene.make_scene(n_cam=4, n_pt=50, mean_track=3.0)
    keep = prob.obs_point < 25
    p2 = sba.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, prob.camera_model, prob.camera_params,
                         prob.points, prob.obs_image[keep], prob.obs_point[keep], prob.obs_xy[keep])
    s2 = p2.counts()
    used_pts = len(np.unique(p2.obs_point))
    used_img = len(np.unique(p2.obs_image))
    assert s2.num_effective_parameters_reduced == 3 * used_pts + 6 * used_img + 4 * used_img


GLOO_WORKER = r'''
import importlib, os, sys, ctypes as C
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
from util import oracle_evaluate
from oracle import loader
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
ora = loader.load(m.capi)
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
prob, _ = m.scene.make_scene(n_cam=6, n_pt=400, mean_track=4.0, seed=11)
pk, ok = prob.shard_masks(2, rank)
ok = ok.astype(bool)
full = oracle_evaluate(ora, m.capi, prob)
# the shard as the library sees it: all images / cameras, the kept points' observations
sub = m.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, prob.camera_model, prob.camera_params, prob.points,
                    prob.obs_image[ok], prob.obs_point[ok], prob.obs_xy[ok], prob.image_pose_constant,
                    prob.image_tvec_constant_mask, prob.camera_param_constant_mask, prob.point_constant)
part = oracle_evaluate(ora, m.capi, sub)
# camera-space gradient of the shard, then the one collective the sharded path needs
g = np.zeros((len(prob.image_camera), 6))
np.add.at(g, sub.obs_image, np.einsum("krj,kr->kj", part["pose"], part["r"]))
t = torch.from_numpy(np.concatenate([[part["cost"]], g.ravel()]))
dist.all_reduce(t)
gf = np.zeros_like(g)
np.add.at(gf, prob.obs_image, np.einsum("krj,kr->kj", full["pose"], full["r"]))
assert abs(t[0].item() - full["cost"]) <= 1e-12 * full["cost"], (t[0].item(), full["cost"])
assert np.allclose(t[1:].numpy().reshape(g.shape), gf, rtol=1e-12, atol=1e-9)
# per-observation residuals are bit-identical whichever rank evaluates them
assert np.array_equal(part["r"], full["r"][ok])
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_world_size_2_gloo_sharded_reduction(sba, oracle, tmp_path):
    """Host-side logic of the N>1 path on CPU: shard by point, evaluate locally, allreduce(sum) the
    camera-space quantities over gloo == the unsharded result."""
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER)
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, str(port), str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


@pytest.mark.gpu
def test_observation_indices_are_checked_on_the_device(sba):
    """sba_create leaves the range check of obs_image / obs_point to its histogram kernels (no host pass over the
    observation list): out-of-range and negative indices are refused with the same error, nothing is built."""
    prob, _ = sba.scene.make_scene(n_cam=6, n_pt=400, mean_track=4.0)
    for field, value in (("obs_point", len(prob.points)), ("obs_point", -1), ("obs_image", len(prob.image_camera)), ("obs_image", -7)):
        arrays = {k: getattr(prob, k).copy() for k in ("obs_image", "obs_point")}
        arrays[field][len(arrays[field]) // 2] = value
        bad = sba.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, prob.camera_model, prob.camera_params,
                              prob.points, arrays["obs_image"], arrays["obs_point"], prob.obs_xy)
        with pytest.raises(sba.capi.SbaError) as e:
            sba.Solver(bad)
        assert e.value.code == sba.capi.ERR_INVALID_ARGUMENT and "out of range" in str(e.value)
    s = sba.Solver(prob)  # and the library is still usable afterwards
    assert s.solve().final_cost > 0
    s.close()


def test_resident_maps_need_a_device(sba):
    """sba_maps_create has no host form: without a device it fails loudly, like sba_create."""
    lib = sba.capi.load()
    if lib.sba_device_count() > 0:
        pytest.skip("a CUDA device is present")
    import numpy as np
    with pytest.raises(sba.capi.SbaError) as e:
        sba.ResidentMaps(np.zeros((2, 4, 4), np.float32))
    assert e.value.code == sba.capi.ERR_NO_DEVICE
    with pytest.raises(ValueError):
        sba.ResidentMaps(np.zeros((4, 4), np.float32))


def test_header_is_plain_c(tmp_path):
    """The boundary is a C ABI: include/sba_b200.h must compile as C99 (what cgo / JNI / a C caller sees), warnings as errors,
    and a C translation unit must link against the library."""
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no C compiler")
    src = tmp_path / "abi.c"
    src.write_text('#include <stdio.h>\n#include "sba_b200.h"\n'
                   'int main(void) { sba_solver_options o; sba_default_solver_options(&o);\n'
                   '  printf("%s %d %d\\n", sba_version(), (int)sizeof(sba_problem_desc), o.max_num_iterations); return 0; }\n')
    pkg = os.path.join(root, "semantic-bundle-adjustment-colmap_b200")
    exe = tmp_path / "abi"
    subprocess.check_call([gcc, "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I" + os.path.join(root, "include"), str(src),
                           "-o", str(exe), "-L" + pkg, "-lsba_b200", "-Wl,-rpath," + pkg])
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and out.stdout.startswith("sba_b200"), out.stdout + out.stderr
