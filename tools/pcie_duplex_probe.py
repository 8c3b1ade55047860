"""Why do duplex PCIe copies run at ~32 + 32 GB/s when each direction alone reaches 48-55 GB/s?  (VERDICT r1
item 8b; the host-to-host figure `e2e` is bound by exactly this.)

Measures, on one GPU, H2D alone / D2H alone / both at once for:
  * pinned memory as torch allocates it (cudaHostAlloc default) vs write-combined for the upload source
    (cudaHostAllocWriteCombined) vs cudaHostRegister of ordinary pages;
  * the process (and its first-touch pages) bound to each NUMA node in turn (set_mempolicy + sched_setaffinity);
  * copy-engine DMA in both directions vs one direction driven by the SMs (a device copy kernel reading /
    writing the pinned host buffer through its UVA mapping) while the copy engine does the other;
  * chunked copies (several cudaMemcpyAsync per direction) vs one large one.
Prints GB/s per direction; wall time by CUDA events on each stream."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

MB = 1 << 20
cudart = ctypes.CDLL("libcudart.so.12")


class HostAlias(object):
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (int(ptr), False), "version": 2,
                                         "strides": None}


def host_alloc(nbytes, flags):
    p = ctypes.c_void_p()
    rc = cudart.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flags))
    if rc != 0:
        raise RuntimeError("cudaHostAlloc(%d, %d) -> %d" % (nbytes, flags, rc))
    return p.value


def memcpy_async(dst, src, nbytes, kind, stream):
    rc = cudart.cudaMemcpyAsync(ctypes.c_void_p(dst), ctypes.c_void_p(src), ctypes.c_size_t(nbytes), ctypes.c_int(kind),
                                ctypes.c_void_p(stream))
    if rc != 0:
        raise RuntimeError("cudaMemcpyAsync -> %d" % rc)


def numa_nodes():
    base = "/sys/devices/system/node"
    try:
        return sorted(int(d[4:]) for d in os.listdir(base) if d.startswith("node") and d[4:].isdigit())
    except OSError:
        return [0]


def node_cpus(node):
    out = set()
    with open("/sys/devices/system/node/node%d/cpulist" % node) as fh:
        for part in fh.read().strip().split(","):
            if not part:
                continue
            a, _, b = part.partition("-")
            out.update(range(int(a), int(b or a) + 1))
    return out


def bind_to_node(node):
    """MPOL_BIND the calling thread's future allocations to `node` and run on its CPUs."""
    libc = ctypes.CDLL(None, use_errno=True)
    mask = ctypes.c_ulong(1 << node)
    rc = libc.syscall(238, 2, ctypes.byref(mask), ctypes.c_ulong(64))         # set_mempolicy(MPOL_BIND)
    cpus = node_cpus(node) & set(os.sched_getaffinity(0))
    if cpus:
        os.sched_setaffinity(0, cpus)
    return rc


def unbind():
    libc = ctypes.CDLL(None, use_errno=True)
    libc.syscall(238, 0, None, ctypes.c_ulong(0))                              # MPOL_DEFAULT


def measure(label, up, down, nbytes, reps=5):
    """up / down: callables (stream) that enqueue one direction's transfer of nbytes, or None."""
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("up", "down", "both"):
        if (mode in ("up", "both") and up is None) or (mode in ("down", "both") and down is None):
            continue
        best = None
        for _ in range(reps):
            torch.cuda.synchronize()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            if mode in ("up", "both"):
                ev[0].record(s_up)
                up(s_up)
                ev[1].record(s_up)
            if mode in ("down", "both"):
                ev[2].record(s_dn)
                down(s_dn)
                ev[3].record(s_dn)
            torch.cuda.synchronize()
            t_up = ev[0].elapsed_time(ev[1]) if mode in ("up", "both") else None
            t_dn = ev[2].elapsed_time(ev[3]) if mode in ("down", "both") else None
            tot = max(t for t in (t_up, t_dn) if t is not None)
            if best is None or tot < best[0]:
                best = (tot, t_up, t_dn)
        res[mode] = best
    gb = nbytes / 1e6

    def rate(t):
        return "  -  " if t is None else "%5.1f" % (gb / t)
    line = "%-64s" % label
    for mode in ("up", "down", "both"):
        if mode in res:
            _, a, b = res[mode]
            line += " | %s: H2D %s D2H %s" % (mode, rate(a), rate(b))
    print(line, flush=True)


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    nbytes = 512 * MB
    n = nbytes // 4
    d_up = torch.empty(n, dtype=torch.float32, device=dev)
    d_dn = torch.ones(n, dtype=torch.float32, device=dev)
    print("NUMA nodes: %s; cpus visible: %d" % (numa_nodes(), len(os.sched_getaffinity(0))))
    try:
        import subprocess
        print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
        bdf = torch.cuda.get_device_properties(0).pci_bus_id if hasattr(torch.cuda.get_device_properties(0), "pci_bus_id") else None
        print("GPU0 pci:", bdf)
        out = subprocess.run("for d in /sys/bus/pci/devices/*; do if [ \"$(cat $d/class)\" = 0x030200 ]; then echo $d $(cat $d/numa_node) "
                             "$(cat $d/current_link_speed) x$(cat $d/current_link_width); fi; done", shell=True, capture_output=True, text=True)
        print(out.stdout)
    except Exception as exc:
        print("topology probe failed:", exc)

    # 1. torch pinned memory, one large copy per direction
    h_up = torch.ones(n, dtype=torch.float32).pin_memory()
    h_dn = torch.empty(n, dtype=torch.float32).pin_memory()

    def dma_up(src_ptr):
        return lambda s: memcpy_async(d_up.data_ptr(), src_ptr, nbytes, 1, s.cuda_stream)

    def dma_dn(dst_ptr):
        return lambda s: memcpy_async(dst_ptr, d_dn.data_ptr(), nbytes, 2, s.cuda_stream)

    measure("torch pinned, one 512 MB copy per direction", dma_up(h_up.data_ptr()), dma_dn(h_dn.data_ptr()), nbytes)

    # 2. chunked
    for chunk_mb in (4, 32):
        c = chunk_mb * MB

        def up_chunks(s, c=c):
            for o in range(0, nbytes, c):
                memcpy_async(d_up.data_ptr() + o, h_up.data_ptr() + o, c, 1, s.cuda_stream)

        def dn_chunks(s, c=c):
            for o in range(0, nbytes, c):
                memcpy_async(h_dn.data_ptr() + o, d_dn.data_ptr() + o, c, 2, s.cuda_stream)
        measure("torch pinned, %d MB chunks" % chunk_mb, up_chunks, dn_chunks, nbytes)

    # 3. write-combined upload source / mapped
    for flags, name in ((0, "cudaHostAlloc default"), (4, "write-combined"), (1, "portable"), (2, "mapped")):
        try:
            p_up = host_alloc(nbytes, flags)
            p_dn = host_alloc(nbytes, 0)
            ctypes.memset(p_up, 1, nbytes)
            ctypes.memset(p_dn, 0, nbytes)
            measure("cudaHostAlloc(%s) source, default destination" % name, dma_up(p_up), dma_dn(p_dn), nbytes)
            cudart.cudaFreeHost(ctypes.c_void_p(p_up))
            cudart.cudaFreeHost(ctypes.c_void_p(p_dn))
        except Exception as exc:
            print("flags", flags, "failed:", exc)

    # 4. SM-driven direction: device copy kernels through the UVA mapping of the pinned buffers
    a_up = torch.as_tensor(HostAlias(h_up.data_ptr(), n), device=dev)
    a_dn = torch.as_tensor(HostAlias(h_dn.data_ptr(), n), device=dev)

    def sm_up(s):
        with torch.cuda.stream(s):
            d_up.copy_(a_up)

    def sm_dn(s):
        with torch.cuda.stream(s):
            a_dn.copy_(d_dn)
    try:
        measure("SM copy kernel up (reads host), DMA down", sm_up, dma_dn(h_dn.data_ptr()), nbytes)
        measure("DMA up, SM copy kernel down (writes host)", dma_up(h_up.data_ptr()), sm_dn, nbytes)
        measure("SM copy kernels both ways", sm_up, sm_dn, nbytes)
    except Exception as exc:
        print("SM-driven copies failed:", exc)

    # 5. NUMA placement of the pinned pages
    nodes = numa_nodes()
    if len(nodes) > 1:
        for node in nodes:
            try:
                all_cpus = os.sched_getaffinity(0)
                rc = bind_to_node(node)
                p_up = host_alloc(nbytes, 0)
                p_dn = host_alloc(nbytes, 0)
                ctypes.memset(p_up, 1, nbytes)
                ctypes.memset(p_dn, 0, nbytes)
                measure("pinned pages + thread on NUMA node %d (set_mempolicy rc %d)" % (node, rc), dma_up(p_up), dma_dn(p_dn), nbytes)
                cudart.cudaFreeHost(ctypes.c_void_p(p_up))
                cudart.cudaFreeHost(ctypes.c_void_p(p_dn))
                unbind()
                os.sched_setaffinity(0, all_cpus)
            except Exception as exc:
                print("node", node, "failed:", exc)
    else:
        print("single NUMA node: placement is not a variable on this box")

    # 6. the pipeline's own strided 2-D copies (128-float slabs of a 602-float matrix), duplex
    from sgc_b200 import _ffi
    rows, f, ld, w = 232965, 602, 608, 128
    hx = torch.ones((rows, f), dtype=torch.float32).pin_memory()
    hy = torch.empty((rows, f), dtype=torch.float32).pin_memory()
    dx = torch.empty((rows, ld), dtype=torch.float32, device=dev)
    dy = torch.ones((rows, ld), dtype=torch.float32, device=dev)
    slab_bytes = rows * w * 4 * 4

    def up2d(s):
        for c0 in range(0, 4 * w, w):
            _ffi.call("sgc_copy_2d", dx.data_ptr() + 4 * c0, 4 * ld, hx.data_ptr() + 4 * c0, 4 * f, 4 * w, rows, s.cuda_stream)

    def dn2d(s):
        for c0 in range(0, 4 * w, w):
            _ffi.call("sgc_copy_2d", hy.data_ptr() + 4 * c0, 4 * f, dy.data_ptr() + 4 * c0, 4 * ld, 4 * w, rows, s.cuda_stream)
    measure("strided 2-D slab copies (4 x 128 of 602 columns, 233 k rows)", up2d, dn2d, slab_bytes)

    def upc(s):
        memcpy_async(dx.data_ptr(), hx.data_ptr(), slab_bytes, 1, s.cuda_stream)

    def dnc(s):
        memcpy_async(hy.data_ptr(), dy.data_ptr(), slab_bytes, 2, s.cuda_stream)
    measure("same bytes, contiguous", upc, dnc, slab_bytes)


if __name__ == "__main__":
    main()
