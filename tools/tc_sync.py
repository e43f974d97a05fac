import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import resolution_pde_b200 as R
R.set_precision(sys.argv[1])
for (M, N, K) in [(1048576, 256, 32), (1048576, 256, 64), (1048576, 256, 256), (262144, 256, 1024)]:
    a = torch.randn(M, K, device="cuda"); b = torch.randn(N, K, device="cuda")
    for _ in range(2): R.ops.matmul(a, b, False, True)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5): R.ops.matmul(a, b, False, True)
    e.record(); torch.cuda.synchronize()
    print(json.dumps({"skip": os.environ.get("RPDE_TC_SKIP"), "prec": sys.argv[1], "shape": [M, N, K], "ms": round(s.elapsed_time(e) / 5, 4)}), flush=True)
