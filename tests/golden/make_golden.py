#!/usr/bin/env python
"""Writes the committed fixtures of tests/golden/.  Run from the repo root:  python tests/golden/make_golden.py

  reference_known_answers.json  every value the REFERENCE's own tests pin for this path, transcribed with file:line
                                (the reference is C++ + Ceres/Eigen/Boost and cannot be compiled or imported in this
                                 container, so no output of the reference itself can be generated -- DESIGN.md section 7)
  oracle_scenes.npz             outputs of the CPU restatement (oracle/) on three small seeded scenes: residuals,
                                Jacobians, cost, and the 10-iteration LM trace.  These travel to the GPU box, where the CUDA
                                path is compared with them (tests/test_golden_fixtures.py); on CPU they pin the oracle
                                against accidental change.
"""
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

KNOWN = {
    "_source": "/root/reference (read-only); values transcribed, not computed",
    "cost_functions": {  # src/base/cost_functions_test.cc:41-69 (variable pose), 71-99 (constant pose), 101-132 (rig)
        "model": "SIMPLE_PINHOLE", "qvec": [1, 0, 0, 0], "tvec": [0, 0, 0], "obs": [0, 0],
        "rig": {"rig_qvec": [1, 0, 0, 0], "rig_tvec": [0, 0, -1], "rel_qvec": [1, 0, 0, 0], "rel_tvec": [0, 0, 1]},
        "cases": [{"X": [0, 0, 1], "params": [1, 0, 0], "residual": [0, 0]},
                  {"X": [0, 1, 1], "params": [1, 0, 0], "residual": [0, 1]},
                  {"X": [0, 1, 1], "params": [2, 0, 0], "residual": [0, 2]},
                  {"X": [-1, 1, 1], "params": [2, 0, 0], "residual": [-2, 2]}]},
    "camera_models": {  # src/base/camera_models_test.cc:133-218; round trips :41-76 on u,v in [-0.5,0.5] step 0.1
        "round_trip_tolerance": 1e-6,
        "params": [[0, [655.123, 386.123, 511.123]], [1, [651.123, 655.123, 386.123, 511.123]],
                   [2, [651.123, 386.123, 511.123, 0]], [2, [651.123, 386.123, 511.123, 0.1]],
                   [3, [651.123, 386.123, 511.123, 0, 0]], [3, [651.123, 386.123, 511.123, 0.1, 0]],
                   [3, [651.123, 386.123, 511.123, 0, 0.05]], [3, [651.123, 386.123, 511.123, 0.05, 0.03]],
                   [4, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001]],
                   [5, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001]],
                   [6, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001, 0.001, 0.02, -0.02, 0.001]],
                   [7, [651.123, 655.123, 386.123, 511.123, 0]], [7, [651.123, 655.123, 386.123, 511.123, 0.9]],
                   [7, [651.123, 655.123, 386.123, 511.123, 1e-6]], [7, [651.123, 655.123, 386.123, 511.123, 1e-2]],
                   [8, [651.123, 386.123, 511.123, 0]], [8, [651.123, 386.123, 511.123, 0.1]],
                   [9, [651.123, 386.123, 511.123, 0, 0]], [9, [651.123, 386.123, 511.123, 0, 0.1]],
                   [9, [651.123, 386.123, 511.123, 0, 0.05]], [9, [651.123, 386.123, 511.123, 0, 0.03]],
                   [10, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001, 0.001, 0.02, -0.02, 0.001]]]},
    "bundle_adjustment_counts": {  # src/optim/bundle_adjustment_test.cc (num_residuals_reduced, num_effective_parameters_reduced)
        "TwoView:210-246": [400, 309], "TwoViewConstantCamera:248-284": [400, 302],
        "PartiallyContainedTracks:286-330": [400, 7], "PartiallyContainedTracksForceToOptimizePoint:332-388": [402, 10],
        "ConstantPoints:390-435": [400, 298], "VariableImage:437-477": [600, 317],
        "ConstantFocalLength:479-529": [400, 307], "VariablePrincipalPoint:531-593": [400, 313],
        "ConstantExtraParam:595-645": [400, 307]},
    "bilinear": {  # src/util/bitmap_test.cc:277-288: 11x11 grey bitmap, pixel (5,5) = 255, rest 0
        "width": 11, "height": 11, "set": [[5, 5, 255]],
        "cases": [[5, 5, 255.0], [5.5, 5, 127.5], [5.5, 5.5, 63.75]]},
}

SCENES = {  # name -> (make_scene kwargs, loss type, loss scale, linear solver)
    "radial_dense": (dict(n_cam=6, n_pt=240, mean_track=4.0, model="SIMPLE_RADIAL", seed=11), 0, 1.0, 0),
    "pinhole_const_intr_iterative": (dict(n_cam=9, n_pt=300, mean_track=4.5, model="SIMPLE_PINHOLE", refine_intr=False,
                                          width=640, height=480, seed=12), 0, 1.0, 2),
    "radial_cauchy_dense": (dict(n_cam=5, n_pt=200, mean_track=3.5, model="SIMPLE_RADIAL", seed=13), 2, 1.5, 0),
}


def main():
    sba = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
    from oracle import loader
    from util import oracle_evaluate, oracle_solve
    ora = loader.load(sba.capi)
    with open(os.path.join(HERE, "reference_known_answers.json"), "w") as f:
        json.dump(KNOWN, f, indent=1)
    out = {}
    for name, (kw, loss, scale, solver) in SCENES.items():
        prob, _ = sba.scene.make_scene(**kw)
        ev = oracle_evaluate(ora, sba.capi, prob, loss_type=loss, loss_scale=scale, threads=1)
        o = sba.capi.default_solver_options()
        o.max_num_iterations, o.linear_solver_type = 10, solver
        o.loss_function_type, o.loss_function_scale = loss, scale
        rc, s, trace, params = oracle_solve(ora, sba.capi, prob, o, threads=1)
        assert rc == 0
        out[name + "/kwargs"] = np.frombuffer(json.dumps(kw).encode(), np.uint8)
        out[name + "/solver"] = np.array([loss, scale, solver], np.float64)
        out[name + "/obs_image"], out[name + "/obs_point"] = prob.obs_image, prob.obs_point
        for k in ("r", "pose", "intr", "point"):
            out[name + "/" + k] = ev[k]
        out[name + "/cost"] = np.array([ev["cost"], s.initial_cost, s.final_cost])
        out[name + "/trace_cost"] = np.array([t.cost for t in trace])
        out[name + "/trace_ok"] = np.array([t.step_is_successful for t in trace], np.int8)
        out[name + "/trace_lin_it"] = np.array([t.linear_solver_iterations for t in trace], np.int32)
        out[name + "/points"] = params[3]
        print(name, prob.num_observations, "obs, cost", s.initial_cost, "->", s.final_cost)
    np.savez_compressed(os.path.join(HERE, "oracle_scenes.npz"), **out)


if __name__ == "__main__":
    main()
