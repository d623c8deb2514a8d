"""GPU helper: TFLOP/s of the tcgen05 GEMM variants on the shapes of the TAN train step against cuBLAS (torch.matmul)
on the same shapes.  Both sides are timed as a CUDA-graph replay of REPS launches (CUDA events around the replay), so
the host cost of a launch (ctypes + five cuTensorMapEncodeTiled calls here, the dispatcher there) is not in the number:
the train step replays a graph too.  GRAPH=0 times eager launches instead."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from exoground_b200 import ops
from exoground_b200.ops import lib as L

K = ops.get()
SHAPES = [  # (M, N, K, tA, tB, epilogue, bias, out_f32, tag)
    (8192, 1536, 512, 0, 0, L.EPI_NONE, 1, 0, "qkv"),
    (8192, 512, 512, 0, 0, L.EPI_NONE, 1, 0, "out_proj"),
    (12288, 2048, 512, 0, 0, L.EPI_QGELU, 1, 0, "c_fc+qgelu"),
    (12288, 512, 2048, 0, 0, L.EPI_NONE, 1, 0, "c_proj"),
    (12288, 2048, 512, 0, 1, L.EPI_NONE, 0, 0, "du (dx @ Wp)"),
    (12288, 2048, 512, 0, 1, L.EPI_QGELU_BWD, 0, 0, "du + qgelu_bwd"),
    (2048, 512, 12288, 1, 1, L.EPI_NONE, 0, 1, "dW_fc splitK"),
    (512, 512, 8192, 1, 1, L.EPI_NONE, 0, 1, "dW_out splitK"),
    (8192, 512, 4096, 0, 0, L.EPI_NONE, 0, 0, "video_pre_proj"),
    (8192, 4096, 512, 0, 0, L.EPI_NONE, 0, 0, "sim (1 stage)"),
    (8192, 8192, 8192, 0, 0, L.EPI_NONE, 0, 0, "square 8192"),
]
impls = [("tc2", L.GEMM_TC2)]
SHAPES += [
    (8192, 2048, 512, 0, 0, L.EPI_QGELU, 1, 0, "c_fc 8192"),
    (8192, 512, 2048, 0, 0, L.EPI_NONE, 1, 0, "c_proj 8192"),
    (12288, 512, 2048, 0, 1, L.EPI_NONE, 0, 0, "dh = du @ Wfc"),
    (12288, 512, 1536, 0, 1, L.EPI_NONE, 0, 0, "dxn = dqkv @ Win"),
    (512, 2048, 12288, 1, 1, L.EPI_NONE, 0, 1, "dW_proj splitK"),
    (1536, 512, 12288, 1, 1, L.EPI_NONE, 0, 1, "dW_in splitK"),
    (12288, 512, 512, 0, 0, L.EPI_NONE, 1, 0, "out_proj 12288"),
]
if os.environ.get("ONLY"):
    SHAPES = [x for x in SHAPES if any(k in x[8] for k in os.environ["ONLY"].split(","))]
PRE = bool(os.environ.get("PRE"))      # c_fc also writes the pre-activation (training forward)
NOCUBLAS = bool(__import__("os").environ.get("NOCUBLAS"))
for (M, N, Kd, tA, tB, epi, bias, f32, tag) in SHAPES:
    A = torch.randn((Kd, M) if tA else (M, Kd), device="cuda").bfloat16()
    B = torch.randn((Kd, N) if tB else (N, Kd), device="cuda").bfloat16()
    bs = torch.randn(N, device="cuda") if bias else None
    out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    aux = torch.randn(M, N, device="cuda").bfloat16() if epi == L.EPI_QGELU_BWD else None
    line = f"{tag:16s} M{M} N{N} K{Kd}: "
    ROT = int(os.environ.get("ROTATE", "0"))
    if ROT:      # cycle through ROT sets of activations / outputs so every call sees L2-cold tensors (as in the step)
        As = [A.clone() for _ in range(ROT)]
        outs = [torch.empty_like(out) for _ in range(ROT)]
        auxs = [aux.clone() if aux is not None else None for _ in range(ROT)]
        for _ in range(3):
            K.gemm(As[0], B, transA=bool(tA), transB=bool(tB), bias=bs, epilogue=epi, out=outs[0], aux=auxs[0],
                   want_preact=PRE and epi == L.EPI_QGELU)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        reps = 24
        for i in range(reps):
            K.gemm(As[i % ROT], B, transA=bool(tA), transB=bool(tB), bias=bs, epilogue=epi, out=outs[i % ROT],
                   aux=auxs[i % ROT], want_preact=PRE and epi == L.EPI_QGELU)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(line + f"rotating x{ROT}: {ms*1e3:8.1f} us {2*M*N*Kd/ms/1e9:7.1f} TF/s", flush=True)
        continue
    def timed(fn, reps=20):
        """us per call of fn(): REPS calls captured into one CUDA graph, 3 replays, best."""
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if os.environ.get("GRAPH", "1") == "0":
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record(); torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps * 1e3
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for _ in range(reps):
                    fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / reps * 1e3)
        return best

    t_own = None
    for name, impl in impls:
        try:
            us = timed(lambda: K.gemm(A, B, transA=bool(tA), transB=bool(tB), bias=bs, epilogue=epi, out=out, impl=impl,
                                      aux=aux, want_preact=PRE and epi == L.EPI_QGELU))
            t_own = us
            line += f"{name} {us:8.1f} us {2*M*N*Kd/us/1e6:7.1f} TF/s | "
        except Exception as ex:
            line += f"{name} ERR {str(ex)[:60]} | "
    if NOCUBLAS:
        print(line, flush=True)
        continue
    Ab = (A.t() if tA else A).contiguous(); Bb = (B if tB else B.t()).contiguous()
    outc = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    us = timed(lambda: torch.matmul(Ab, Bb, out=outc))
    line += f"cublas {us:8.1f} us {2*M*N*Kd/us/1e6:7.1f} TF/s"
    if t_own is not None:
        line += f" | own/cublas time {t_own/us:5.2f}" + ("  WIN" if t_own <= us else "")
    print(line, flush=True)
