"""Library FP64 numbers on the box (cuBLAS DGEMM / cuSOLVER batched potrf through torch).

Context only: these are the *library* figures our hand-written kernels are compared against in
DESIGN.md; nothing in the product path calls them.  Prints one JSON object.
"""
import json
import time

import torch


def _time(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    out = {"gpu": torch.cuda.get_device_name(0)}
    dev = "cuda:0"
    for n in (2048, 4096, 8192):
        a = torch.randn(n, n, dtype=torch.float64, device=dev)
        b = torch.randn(n, n, dtype=torch.float64, device=dev)
        ms = _time(lambda: torch.matmul(a, b))
        out[f"cublas_dgemm_{n}_tflops"] = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    # sustained 3 s
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    t0 = time.time()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    k = 0
    while time.time() - t0 < 3.0:
        for _ in range(4):
            torch.matmul(a, b)
        k += 4
        torch.cuda.synchronize()
    e1.record()
    e1.synchronize()
    out["cublas_dgemm_8192_sustained3s_tflops"] = 2.0 * n ** 3 * k / (e0.elapsed_time(e1) * 1e-3) / 1e12
    del a, b
    # batched Cholesky + inverse through the library (what TF/torch users get)
    for (bsz, n) in ((64, 1024), (32, 2048)):
        x = torch.randn(bsz, n, n, dtype=torch.float64, device=dev)
        k = x @ x.transpose(1, 2) + n * torch.eye(n, dtype=torch.float64, device=dev)
        del x
        ms = _time(lambda: torch.linalg.cholesky(k), reps=3, warm=1)
        out[f"lib_potrf_b{bsz}_n{n}_tflops"] = bsz * n ** 3 / 3 / (ms * 1e-3) / 1e12
        L = torch.linalg.cholesky(k)
        ms = _time(lambda: torch.cholesky_inverse(L), reps=3, warm=1)
        out[f"lib_potri_b{bsz}_n{n}_tflops"] = bsz * 2 * n ** 3 / 3 / (ms * 1e-3) / 1e12
        del k, L
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
