import sys, os, faulthandler, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ocsmesh_b200 as ob
gold = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests/golden/reference_golden.npz"))
lines = [gold[f"feat_line_{k}"] for k in range(int(gold["feat_n_lines"]))]
n = int(sys.argv[1])
r = ob.Raster.from_numpy(gold["dem"], gold["dem_transform"], crs=32618, chunk_size=40, overlap=2)
t0 = time.time()
for i in range(n):
    faulthandler.dump_traceback_later(20, exit=True)
    hf = ob.Hfun(r, hmin=30.0, hmax=None)
    hf.add_feature(lines, 0.01, None)
    hf.add_subtidal_flow_limiter(50.0, None, -30.0, 0.0)
    hf.add_gradient_delimiter(None, None)
    ok = np.array_equal(hf.get_values(), gold["hfun_no_hmax_gradient"])
    faulthandler.cancel_dump_traceback_later()
    if not ok: print("MISMATCH at", i, flush=True)
    if i % 20 == 0: print("iter", i, ok, round(time.time() - t0, 2), flush=True)
print("done", n, flush=True)
