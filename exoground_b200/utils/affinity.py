"""Host-side placement for one-process-per-GPU jobs: run the rank's Python thread (and therefore the first touch of its
pinned staging pool) on the CPUs of the NUMA node its GPU hangs off.

Why: every rank copies its batch from pinned host memory each step (TAN configs[1]: 201 MB of fp32 features per step and
GPU).  Eight unbound ranks on a two-socket host allocate their pinned pools wherever the launcher happened to start
them, so half of the H2D traffic crosses the socket interconnect and all of it competes on one memory controller set:
round 1 measured 0.79 end-to-end scaling at 8 GPUs against 0.94 device-timed (profiles/scale_r01f_summary.txt).  The
reference leaves placement to torch's DataLoader workers (train/main.py:404-411, `pin_memory=True`); this is the
one-process-per-GPU equivalent of launching under `numactl --cpunodebind --membind`.
"""
import os


def gpu_cpu_set(index):
    """CPUs NVML reports as local to GPU `index` (empty set when NVML cannot say)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByIndex(int(index))
            n_cpu = os.cpu_count() or 64
            words = (n_cpu + 63) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
            cpus = {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        finally:
            pynvml.nvmlShutdown()
        return cpus
    except Exception:  # noqa: BLE001  (no NVML, no permission, container without topology: stay where we are)
        return set()


def bind_to_gpu(index):
    """Restrict this process to the CPUs local to GPU `index` that the container lets it use.  Call BEFORE allocating
    pinned host memory.  Returns a short description for logs / bench config ("cpus 0-31 (32)" or "unchanged: ...")."""
    if not hasattr(os, "sched_setaffinity"):
        return "unchanged: no sched_setaffinity"
    local = gpu_cpu_set(index)
    if not local:
        return "unchanged: NVML reports no CPU affinity"
    allowed = os.sched_getaffinity(0)
    target = sorted(local & allowed)
    if not target:
        return "unchanged: GPU-local CPUs are outside this process's cpuset"
    if set(target) == set(allowed):
        return f"unchanged: already on the {len(target)} GPU-local CPUs"
    try:
        os.sched_setaffinity(0, target)
    except OSError as e:
        return f"unchanged: {e}"
    return f"cpus {target[0]}-{target[-1]} ({len(target)})"
