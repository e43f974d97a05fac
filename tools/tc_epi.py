import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import resolution_pde_b200 as R
R.set_precision(sys.argv[1])
R.ops.matmul(torch.randn(256, 256, device="cuda"), torch.randn(256, 256, device="cuda"))   # syncs the engine choice
lib = R._lib.load(); st = torch.cuda.current_stream().cuda_stream
T = 1048576
def tm(fn):
    for _ in range(2): fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5): fn()
    e.record(); torch.cuda.synchronize()
    return round(s.elapsed_time(e) / 5, 4)
for (i, o) in [(64, 256), (256, 256)]:
    x = torch.randn(T, i, device="cuda"); w = torch.randn(o, i, device="cuda") / i ** 0.5; b = torch.randn(o, device="cuda")
    y = torch.empty(T, o, device="cuda"); z = torch.empty(T, o, device="cuda")
    r = {"dbg": os.environ.get("RPDE_TC_DBG"), "prec": sys.argv[1], "in_out": [i, o]}
    r["plain"] = tm(lambda: lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), None, T, i, o, 0, 0, st))
    r["gelu"] = tm(lambda: lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), None, T, i, o, 1, 0, st))
    r["gelu_z"] = tm(lambda: lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), z.data_ptr(), T, i, o, 1, 0, st))
    r["gelu_gradz"] = tm(lambda: lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), z.data_ptr(), T, i, o, 0x101, 0, st))
    print(json.dumps(r), flush=True)
# dgrad with the stored act' multiply (+ bias-gradient column sums): dx(T x in) = dy(T x out) . W(out x in) * saved(T x in)
for (i, o) in [(256, 64), (256, 256), (64, 256)]:
    dy = torch.randn(T, o, device="cuda"); w = torch.randn(o, i, device="cuda") / i ** 0.5
    sv = torch.rand(T, i, device="cuda"); dx = torch.empty(T, i, device="cuda"); db = torch.zeros(i, device="cuda")
    r = {"prec": sys.argv[1], "dgrad_in_out": [i, o]}
    r["plain"] = tm(lambda: lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), None, 0, dx.data_ptr(), 0, None, T, i, o, 0, st))
    r["saved"] = tm(lambda: lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), sv.data_ptr(), 3, dx.data_ptr(), 0, None, T, i, o, 0, st))
    r["saved_dbias"] = tm(lambda: lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), sv.data_ptr(), 3, dx.data_ptr(), 0, db.data_ptr(), T, i, o, 0, st))
    print(json.dumps(r), flush=True)
