import sys, time, ctypes as C
sys.path.insert(0, "/root/repo")
import torch
from morphablegraphs_b200 import _lib
lib = _lib.load()
M, d = 1000000, 1500
X = torch.randn((M, d), device="cuda") * 30
def t(fn, n=3):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("colsum double ms", t(lambda: X.double().sum(dim=0), 1))
print("colsum f32->f64 chunks ms", t(lambda: sum(X[a:a+65536].sum(dim=0, dtype=torch.float64) for a in range(0, M, 65536))))
mean = X.mean(dim=0).contiguous()
up = torch.zeros((d, d), dtype=torch.float64, device="cuda")
def syrk():
    _lib.check(lib.mgb_syrk_upper(C.c_void_p(X.data_ptr()), M, d, d, C.c_void_p(mean.data_ptr()), C.c_void_p(up.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
ms = t(syrk)
print("syrk ms", ms, "TFLOP/s (algorithmic 2 n d^2)", 2 * M * d * d / ms / 1e9)
cov = up + up.t()
print("eigh ms", t(lambda: torch.linalg.eigh(cov), 1))
W = torch.randn((d, 34), device="cuda")
print("project ms", t(lambda: [(X[a:a+131072] - mean) @ W for a in range(0, M, 131072)]))
Xc = X[:131072] - mean
print("torch matmul cov chunk ms x8", 8 * t(lambda: (Xc.t() @ Xc)))
