import importlib, sys, os, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sba = importlib.import_module('semantic-bundle-adjustment-colmap_b200')
from oracle import loader
oracle = loader.load(sba.capi)
import test_gpu_semantic as T
capi = sba.capi
q, t, cams, depth, label = T.plane_scene(sba)
blocks = T.pair_blocks(4, 120, 160, 5, depth)
d, keep = T.sem_desc(sba, q, t, cams, depth, label, blocks)
n = d.num_blocks
rg, sg, jg = np.zeros(n), np.zeros(n, np.int32), np.zeros((n, 14))
ro, so, jo = np.zeros(n), np.zeros(n, np.int32), np.zeros((n, 14))
capi.check(capi.load().sba_semantic_evaluate(C.byref(d), capi.ptr(rg, capi._dp), capi.ptr(sg, capi._ip), capi.ptr(jg, capi._dp), None))
oracle.oracle_semantic_evaluate(C.byref(d), 0, capi.ptr(ro, capi._dp), capi.ptr(so, capi._ip), capi.ptr(jo, capi._dp))
print("nan per column", np.isnan(jg).sum(0))
print("rows with nan", np.isnan(jg).any(1).sum(), "of", n)
bad = np.nonzero(~np.isnan(jg).any(1) & (jg != jo).any(1))[0]
print("non-nan mismatching rows", len(bad))
print(jg[0], jo[0])
i = np.nonzero(jo.any(1))[0][:3]
print(jg[i], jo[i])
