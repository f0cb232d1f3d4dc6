"""Microbenchmark of the tcgen05 GEMM tile through hb_gemm_f16 (CUDA events, L2-cold not enforced:
operands of the larger shapes exceed L2).  python tools/gemm_bench.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from hanabi_agents_b200 import _lib
    L = _lib.lib()
    dev = "cuda"
    if len(sys.argv) > 3:
        shapes = [tuple(int(x) for x in sys.argv[1:4])]
    else:
      shapes = [(65536, 512, k) for k in (64, 256, 768, 2048, 4096)] + [(65536, 1280, 512), (65536, 256, 768),
                                                                        (65536, 128, 768), (8192, 8192, 8192)]
    modes = (int(sys.argv[4]),) if len(sys.argv) > 4 else (0, 1, 2)
    for M, N, K in shapes:
        a = torch.randn(M, K, device=dev).half()
        b = (torch.randn(N, K, device=dev) / K ** 0.5).half()
        bias = torch.zeros(N, device=dev)
        for mode in modes:
            if mode == 2 and N % 64:
                continue
            out = torch.empty((M, N), dtype=torch.float16 if mode == 0 else torch.float32, device=dev)
            q = torch.empty((M, N // 64), dtype=torch.float32, device=dev)

            def run():
                _lib.check(L.hb_gemm_f16(a.data_ptr(), K, b.data_ptr(), K, M, N, K, mode, 1.0, bias.data_ptr(), 1,
                                         None if mode == 2 else out.data_ptr(), N, None, 0, q.data_ptr(), 51,
                                         -25.0, 1.0, 1 / 51, _lib.current_stream_ptr()))
            for _ in range(3):
                run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            e0.record()
            for _ in range(reps):
                run()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            print("M=%d N=%d K=%d mode=%d: %.1f us  %.0f TFLOP/s" % (M, N, K, mode, ms * 1e3, 2.0 * M * N * K / ms / 1e9))
        for _ in range(3):
            c = a @ b.t()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            c = a @ b.t()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print("M=%d N=%d K=%d cuBLAS fp16: %.1f us  %.0f TFLOP/s" % (M, N, K, ms * 1e3, 2.0 * M * N * K / ms / 1e9))


if __name__ == "__main__":
    main()
