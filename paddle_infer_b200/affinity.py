"""Host-thread placement for the host-buffer path (ops.HostFrontEnd).

In the zero-copy mode the kernel reads ids / LoD from pinned host memory and writes pooled rows
back to it over PCIe.  Pinned pages land on the NUMA node of the thread that allocates them; on
a two-socket box a buffer on the far socket puts every PCIe transaction on the inter-socket
link as well.  `bind_to_gpu(device)` pins the calling thread (and the threads it starts later)
to the CPUs NVML reports as local to the GPU, so that pinned buffers allocated afterwards are
local too.  Set-up helper: nothing on the data path depends on it."""
import os


def _nvml_handle(device_index: int):
    import pynvml as nv
    nv.nvmlInit()
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        entry = vis.split(",")[device_index].strip()
        if entry.isdigit():
            return nv, nv.nvmlDeviceGetHandleByIndex(int(entry))
        return nv, nv.nvmlDeviceGetHandleByUUID(entry)
    return nv, nv.nvmlDeviceGetHandleByIndex(device_index)


def local_cpus(device_index: int):
    """The CPUs NVML lists as local to the GPU (a sorted list), or None when NVML cannot tell."""
    try:
        nv, h = _nvml_handle(device_index)
        n_cpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1]
        return sorted(c for c in cpus if c < n_cpu) or None
    except Exception:
        return None


def bind_to_gpu(device_index: int):
    """Restricts the calling thread to the GPU-local CPUs.  Returns the previous affinity set
    (hand it to `restore`) or None when nothing was changed."""
    cpus = local_cpus(device_index)
    if not cpus or not hasattr(os, "sched_setaffinity"):
        return None
    before = os.sched_getaffinity(0)
    allowed = before & set(cpus)
    if not allowed or allowed == before:
        return None
    os.sched_setaffinity(0, allowed)
    return before


def restore(previous):
    if previous:
        os.sched_setaffinity(0, previous)
