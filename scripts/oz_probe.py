"""GPU probe of the int8-slice emulated fp64 GEMM (csrc/ozaki_i8.cu): accuracy against torch fp64 and timing."""
import json
import sys

import torch

from gaussian_processes_b200 import native


def err(C, ref, A, B):
    # error relative to the DGEMM-style bound sum_k |a||b|
    bound = A.abs() @ B.abs().T
    return float(((C - ref).abs() / bound).max()), float((C - ref).abs().max() / ref.abs().max())


def main():
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(3)
    out = {}
    for (M, N, K) in [(128, 64, 64), (256, 128, 128), (300, 70, 100), (1000, 500, 1600)]:
        A = torch.randn(M, K, generator=g, dtype=torch.float64, device=dev)
        B = torch.randn(N, K, generator=g, dtype=torch.float64, device=dev) * 3.0
        ref = A @ B.T
        for S in (8, 7, 6):
            C = native.gemm_i8emu(A, B, S)
            torch.cuda.synchronize()
            out[f"{M}x{N}x{K}_S{S}"] = err(C, ref, A, B)
            print(M, N, K, S, out[f"{M}x{N}x{K}_S{S}"], flush=True)
    # wide dynamic range rows
    A = torch.randn(512, 256, generator=g, dtype=torch.float64, device=dev)
    A *= torch.exp(8 * torch.randn(512, 1, generator=g, dtype=torch.float64, device=dev))
    B = torch.randn(192, 256, generator=g, dtype=torch.float64, device=dev)
    C = native.gemm_i8emu(A, B, 8)
    print("rowscaled", err(C, A @ B.T, A, B), flush=True)
    # contraction over rows (X^T R shape), split-K, ragged sizes
    for (K, M, N) in [(300, 70, 50), (5000, 200, 130), (70000, 160, 100)]:
        At = torch.randn(K, M, generator=g, dtype=torch.float64, device=dev)
        Bt = torch.randn(K, N, generator=g, dtype=torch.float64, device=dev)
        Bt *= torch.exp(3 * torch.randn(1, N, generator=g, dtype=torch.float64, device=dev))
        C = native.gemm_i8emu(At, Bt, 8, transposed=True)
        print("transposed", K, M, N, err(C, At.T @ Bt, At.T, Bt.T), flush=True)
    if len(sys.argv) > 1 and sys.argv[1] == "time":
        Kt, Mt, Nt = 65536, 1600, 500
        At = torch.randn(Kt, Mt, generator=g, dtype=torch.float64, device=dev)
        Bt = torch.randn(Kt, Nt, generator=g, dtype=torch.float64, device=dev)
        for _ in range(2):
            native.gemm_i8emu(At, Bt, 8, transposed=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            native.gemm_i8emu(At, Bt, 8, transposed=True)
        e1.record()
        torch.cuda.synchronize()
        print(f"X^T R shape S=8: {e0.elapsed_time(e1) / 5:.3f} ms per call incl. slicing both operands", flush=True)
        del At, Bt
        M, N, K = 65536, 500, 1600
        A = torch.randn(M, K, generator=g, dtype=torch.float64, device=dev)
        B = torch.randn(N, K, generator=g, dtype=torch.float64, device=dev)
        for S in (8, 7):
            for _ in range(2):
                native.gemm_i8emu(A, B, S)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                native.gemm_i8emu(A, B, S)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            print(f"S={S}: {ms:.3f} ms per call incl. slicing; fp64-equivalent {2 * M * N * K / ms / 1e9:.1f} TFLOP/s",
                  flush=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        native.matmul_nt(A, B)
        e0.record()
        for _ in range(5):
            native.matmul_nt(A, B)
        e1.record()
        torch.cuda.synchronize()
        print(f"DMMA: {e0.elapsed_time(e1) / 5:.3f} ms", flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
