#!/usr/bin/env python
"""bench.py — GB/s of model bytes chunked + chunk-hashed + file-hashed, bit-exact, on N B200s.

    python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K --warmup W

Headline workload (BASELINE.json configs[1]): the Llama-3-8B bf16 checkpoint shape — four safetensors-shaped
shards of 4.98 / 5.00 / 4.92 / 1.17 GB (16.06 GB) of counter-based splitmix64 bytes.
At N > 1 every rank holds its own copy of that shape with different seeds (file-partitioned, weak scaling: no
data-path collective; only the 32-byte file hashes are gathered).

One step = one pass of the whole hot path (K1 gear candidates -> K2 boundary resolution -> K3 keyed BLAKE3
chunk hashes -> K4 merkle + file hash) over all shards of the rank.
  value : inputs already resident in HBM, timed with CUDA events on the stream the library launches on
  e2e   : the same call through the C ABI with the shards in pinned HOST memory (H2D inside the timed region)
Inputs (16 GB per rank) are far larger than the 126 MB L2, so no explicit L2 flush is needed.

Parity gate: before anything is timed, the device-resident bytes of the headline workload are copied to tmpfs
and hashed by hf_xet (the compiled xet-core); `value` is only printed when every file hash agrees.

Further records on the same JSON line:
  configs.c1 / configs.c5      (N = 1)  BASELINE configs 1 and 5, device-resident, with their own parity flags
  config3                      (N > 1)  BASELINE config 3: 30 shards (138 GB) LPT-partitioned by file, strong scaling
  config4                      (N = 8)  BASELINE config 4: 163 shards (689 GB) laid end to end and cut into N byte
                                        ranges; straddling shards use the halo + carried-cut protocol over NCCL P2P
  h2d_ceiling                  (N > 1)  all ranks copying their pinned shards at once, no kernels

`--impl reference` times the reference's own CPU arithmetic: hf_xet.hash_files — the official compiled build of
huggingface/xet-core (the crate /root/reference/pkg/xet pins in Cargo.lock; its Rust sources are not vendored and
cannot be compiled here) — on the SAME 16.06 GB shard set, generated on the host into tmpfs.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "GB/s model bytes chunked+hashed bit-exact"
# real Llama-3-8B safetensors shard sizes (model-0000{1..4}-of-00004.safetensors)
LLAMA3_8B_SHARDS = [4_976_698_672, 4_999_802_720, 4_915_916_176, 1_168_138_808]
SEED_BASE = 10
# BASELINE config 3: Llama-3-70B bf16 shape, 30 shards (the last one short, as in real checkpoints), seeds 100..129
C3_SIZES = [4_700_000_000 + 1_000_003 * (i % 5) for i in range(29)] + [2_100_000_000]
C3_SEED = 100
# BASELINE config 4: DeepSeek-V3 fp8 shape, 163 shards of ~4.23 GB (not all equal), seeds 1000..1162
C4_SIZES = [4_230_000_000 + 4096 * (i % 7) + (i % 3) for i in range(163)]
C4_SEED = 1000
# lane-instructions per input byte of the two hot kernels, from the committed ncu capture (profiles/*_ncu_summary.json)
NCU_SUMMARY = os.path.join(ROOT, "profiles", "r2_ncu_summary.json")
NCU_SUMMARY_OLD = os.path.join(ROOT, "profiles", "r1_ncu_summary.json")


# ------------------------------------------------------------------------------------- shard contents
def safetensors_header(payload_bytes: int) -> bytes:
    """8-byte LE header length + JSON header padded to 8 bytes (pkg/hfutil/modelconfig/safetensors.go:78-100)."""
    n_elems = payload_bytes // 2
    hdr = json.dumps({"__metadata__": {"format": "pt"},
                      "t0": {"dtype": "BF16", "shape": [n_elems], "data_offsets": [0, n_elems * 2]}},
                     separators=(",", ":")).encode()
    hdr += b" " * (-len(hdr) % 8)
    return len(hdr).to_bytes(8, "little") + hdr


def shard_header(size: int) -> bytes:
    return safetensors_header(size - 256)


def shard_layout(sizes, scale: float):
    sizes = [max(1 << 20, int(s * scale)) for s in sizes]
    offsets = [0]
    for s in sizes:
        offsets.append(offsets[-1] + s)
    return sizes, offsets


def fill_file_range(hasher, buf, dst_off: int, size: int, seed: int, begin: int = 0, end: int = None):
    """Device bytes [begin, end) of the safetensors-shaped shard (size, seed) -> buf[dst_off ...]: header bytes
    through a small H2D copy, payload generated on the device (K0)."""
    import torch
    end = size if end is None else end
    hdr = shard_header(size)
    if begin < len(hdr) and end > begin:
        part = hdr[begin:min(end, len(hdr))]
        buf[dst_off:dst_off + len(part)].copy_(torch.frombuffer(bytearray(part), dtype=torch.uint8))
    lo = max(begin, len(hdr))
    if end > lo:
        hasher.fill_splitmix(buf.data_ptr() + dst_off + (lo - begin), seed, lo - len(hdr), end - lo)


def pwrite_all(fd, arr, off: int):
    mv = memoryview(arr).cast("B")
    done = 0
    while done < len(mv):
        done += os.pwrite(fd, mv[done:], off + done)


def write_shard_file(path: str, size: int, seed: int, begin: int = 0, end: int = None, slab: int = 64 << 20):
    """The same bytes written on the HOST (oracle/xet_oracle.c splitmix fill) into `path` at offset `begin`."""
    import xet_oracle
    end = size if end is None else end
    hdr = shard_header(size)
    fd = os.open(path, os.O_CREAT | os.O_WRONLY, 0o600)
    try:
        if begin < len(hdr) and end > begin:
            os.pwrite(fd, hdr[begin:min(end, len(hdr))], begin)
        pos = max(begin, len(hdr))
        while pos < end:
            m = min(slab, end - pos)
            pwrite_all(fd, xet_oracle.splitmix_bytes(seed, m, pos - len(hdr)), pos)
            pos += m
    finally:
        os.close(fd)


def run_threads(jobs, n_threads):
    """Run callables on a small thread pool (the C fill and file writes release the GIL)."""
    jobs = list(jobs)
    lock, errs = threading.Lock(), []

    def work():
        while True:
            with lock:
                if not jobs:
                    return
                j = jobs.pop()
            try:
                j()
            except Exception as e:  # noqa: BLE001
                errs.append(e)

    ts = [threading.Thread(target=work) for _ in range(max(1, min(n_threads, len(jobs))))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    if errs:
        raise errs[0]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.idx)],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], 0, set(), []
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = max(smax, float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        sm.sort()
        # median of the samples under load (upper half: the sampler also sees the idle edges)
        load = sm[len(sm) // 2:] if sm else []
        med = load[len(load) // 2] if load else None
        return {"sm_mhz": med, "sm_max_mhz": smax or None, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(power) if power else None}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def tmp_root():
    for d in ("/dev/shm", tempfile.gettempdir()):
        if os.path.isdir(d) and os.access(d, os.W_OK):
            return d
    return "."


def tmp_free_bytes():
    try:
        return shutil.disk_usage(tmp_root()).free
    except OSError:
        return 0


# ------------------------------------------------------------------------------------------- CPU side
def cpu_hash_files(paths):
    """The reference arithmetic on the host: hf_xet.hash_files (compiled xet-core) -> (seconds, hex hashes, kind).
    Falls back to the repo's C oracle, one thread per file, if hf_xet cannot be imported."""
    try:
        import hf_xet
        t0 = time.perf_counter()
        infos = hf_xet.hash_files(list(paths))
        return time.perf_counter() - t0, [i.hash for i in infos], "reference"
    except ImportError:
        import numpy as np
        import xet_oracle
        out = [None] * len(paths)

        def one(i):
            a = np.fromfile(paths[i], dtype=np.uint8)
            out[i] = xet_oracle.to_hex(xet_oracle.hash_file(a).file_hash)

        t0 = time.perf_counter()
        run_threads([lambda i=i: one(i) for i in range(len(paths))], os.cpu_count() or 4)
        return time.perf_counter() - t0, out, "port"


def run_reference(args):
    """Reference arm: hf_xet.hash_files (compiled xet-core) on the host cores over the full shard set; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sizes, _ = shard_layout(LLAMA3_8B_SHARDS, args.scale)
    tmpdir = tempfile.mkdtemp(prefix="xetbench-ref-", dir=tmp_root())
    try:
        paths = [os.path.join(tmpdir, f"shard-{i}.safetensors") for i in range(len(sizes))]
        # four shards, each written by several threads in disjoint ranges
        jobs = []
        for i, (p, n) in enumerate(zip(paths, sizes)):
            step = max(64 << 20, -(-n // 8))
            for b in range(0, n, step):
                jobs.append(lambda p=p, n=n, i=i, b=b, step=step: write_shard_file(p, n, SEED_BASE + i, b, min(n, b + step)))
        run_threads(jobs, os.cpu_count() or 8)
        total = sum(sizes)
        kind = "reference"
        for _ in range(max(1, min(args.warmup, 3))):
            _, _, kind = cpu_hash_files(paths)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            _, hexes, kind = cpu_hash_files(paths)
        dt = time.perf_counter() - t0
    finally:
        shutil.rmtree(tmpdir, ignore_errors=True)
    value = total * args.steps / dt / 1e9
    what = "hf_xet 1.5.0 hash_files (compiled huggingface/xet-core)" if kind == "reference" else "oracle/xet_oracle.c, one thread per file"
    sample = (f"{what} on the full Llama-3-8B-shaped shard set ({len(sizes)} files, {total / 1e9:.2f} GB, "
              f"the bytes of the GPU arm's rank 0) in tmpfs, one call per step")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "Llama-3-8B bf16 checkpoint shape: 4 safetensors-shaped shards"
                               + ("" if args.scale == 1.0 else f" (SCALED x{args.scale}: not the named config)"),
                   "shard_bytes": sizes, "bytes": total, "file_hashes": hexes},
        "cpu_baseline": {"value": round(value, 4), "unit": "GB/s", "cores": os.cpu_count(), "kind": kind,
                         "sample": sample},
        "e2e": {"value": round(value, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def bind_to_gpu_numa(device_index: int):
    """Best effort: run this rank on the CPU cores NVML reports as local to its GPU, so that the pinned
    host shards (first-touch) and the copy threads sit on the GPU's NUMA node.  Returns the core count."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cores = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        cores = [c for c in cores if c in os.sched_getaffinity(0)]
        if cores:
            os.sched_setaffinity(0, cores)
            return len(cores)
    except Exception:
        pass
    return None


# ------------------------------------------------------------------------------------------- GPU side
class Ctx:
    """What every leg of the GPU arm shares."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.args = args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        self.numa_cores = bind_to_gpu_numa(self.local_rank) if self.world > 1 else None
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
        # the library launches on this stream; torch ops, NCCL and the timing events use it too, so everything a
        # step does is ordered on one stream and bracketed by the events
        self.stream = torch.cuda.Stream()
        torch.cuda.set_stream(self.stream)
        self.tmpdir = None

    def barrier(self):
        import torch
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        import torch
        import torch.distributed as dist
        t = torch.tensor([x], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_true(self, ok: bool) -> bool:
        import torch
        import torch.distributed as dist
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    def shared_tmpdir(self, tag: str) -> str:
        """One tmpfs directory all ranks of this node agree on (rank 0 creates it)."""
        import torch.distributed as dist
        name = [tempfile.mkdtemp(prefix=f"xetbench-{tag}-", dir=tmp_root()) if self.rank == 0 else None]
        if self.world > 1:
            dist.broadcast_object_list(name, src=0)
        return name[0]

    def timed(self, fn, steps: int, warmup: int):
        """warmup untimed steps, then `steps` timed ones between barriers; returns ms per step, max over ranks."""
        import torch
        for _ in range(warmup):
            fn()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for _ in range(steps):
            fn()
        e1.record(self.stream)
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1)) / steps


def device_file_to_tmpfs(buf, off: int, n: int, path: str, file_off: int = 0, slab: int = 256 << 20):
    """Copy device bytes buf[off:off+n] into `path` at `file_off` (what the parity legs hand to the CPU reference)."""
    fd = os.open(path, os.O_CREAT | os.O_WRONLY, 0o600)
    try:
        pos = 0
        while pos < n:
            m = min(slab, n - pos)
            pwrite_all(fd, buf[off + pos: off + pos + m].cpu().numpy(), file_off + pos)
            pos += m
    finally:
        os.close(fd)


def leg_config2(cx: Ctx):
    """Headline: config 2 device-resident (`value`), full-size parity gate, e2e from pinned host memory."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import ome_b200
    args, world, rank = cx.args, cx.world, cx.rank
    sizes, offsets = shard_layout(LLAMA3_8B_SHARDS, args.scale)
    total = offsets[-1]
    hasher = ome_b200.XetHasher(device=cx.local_rank, stream=cx.stream.cuda_stream, stage_timing=True)
    buf = torch.empty(total + 256, dtype=torch.uint8, device="cuda")
    seed0 = SEED_BASE + 100 * rank
    for i, size in enumerate(sizes):
        fill_file_range(hasher, buf, offsets[i], size, seed0 + i)
    torch.cuda.synchronize()
    off_np = np.asarray(offsets, dtype=np.uint64)

    hashes_dev = torch.zeros(len(sizes), 32, dtype=torch.uint8, device="cuda")
    gathered = [torch.zeros_like(hashes_dev) for _ in range(world)] if world > 1 else None

    def step_device():
        cl = hasher.hash_buffers(buf, off_np, memory=ome_b200.hasher.MEM_DEVICE)
        if world > 1:  # the only exchange of the file-partitioned path: 32 bytes per file to every rank
            hashes_dev.copy_(torch.from_numpy(np.frombuffer(b"".join(f.file_hash for f in cl.files),
                                                            dtype=np.uint8).reshape(-1, 32).copy()))
            dist.all_gather(gathered, hashes_dev)
        return cl

    # ---- parity gate at full size, BEFORE any timing: the device-resident bytes -> tmpfs -> hf_xet.
    # Rank 0 checks all of its shards; at N > 1 every other rank checks its smallest shard (the host has one set
    # of cores for all ranks), so every GPU's result is compared with the CPU reference on real bytes.
    cl = step_device()
    gpu_hex = [f.hex for f in cl.files]
    parity = {"full_config": None}
    cpu = None
    if not args.no_cpu:
        which = list(range(len(sizes))) if rank == 0 else [min(range(len(sizes)), key=lambda i: sizes[i])]
        need = sum(sizes[i] for i in which)
        tmpdir = tempfile.mkdtemp(prefix=f"xetbench-r{rank}-", dir=tmp_root())
        try:
            ok_space = tmp_free_bytes() > need * (1.5 if world == 1 else world)
            if not ok_space and rank == 0:
                which = which[-1:]
            paths = [os.path.join(tmpdir, f"shard-{i}.safetensors") for i in which]
            for i, p in zip(which, paths):
                device_file_to_tmpfs(buf, offsets[i], sizes[i], p)
            if rank == 0:
                cpu_hash_files(paths[-1:])  # warm the library
            cx.barrier()  # rank 0's timed CPU run below should not compete with other ranks' tmpfs writes
            dt, ref_hex, kind = cpu_hash_files(paths)
            ok = ref_hex == [gpu_hex[i] for i in which]
            n_bytes = sum(sizes[i] for i in which)
            if rank == 0:
                what = "hf_xet 1.5.0 hash_files (compiled huggingface/xet-core; one call, file-parallel)" if kind == "reference" \
                    else "oracle/xet_oracle.c, one thread per file"
                cpu = {"value": round(n_bytes / dt / 1e9, 4), "unit": "GB/s", "cores": os.cpu_count(), "kind": kind,
                       "sample": f"{what} on {'the full shard set' if len(which) == len(sizes) else 'the smallest shard'} "
                                 f"({len(which)} files, {n_bytes / 1e9:.2f} GB, copied from the device buffer to tmpfs), {dt:.2f} s wall"
                                 + (f", while {world - 1} other ranks hash one shard each on the same cores" if world > 1 else "")}
                if world == 1:
                    # the same files through the repo's own file entry point (paths in, hashes out)
                    g = hasher.hash_files(paths)
                    t0 = time.perf_counter()
                    g = hasher.hash_files(paths)
                    dt_gpu = time.perf_counter() - t0
                    ok = ok and [f.hex for f in g.files] == ref_hex
                    parity["same_files_through_xetgpu_hash_files"] = {
                        "value": round(n_bytes / dt_gpu / 1e9, 2), "unit": "GB/s", "equal": [f.hex for f in g.files] == ref_hex,
                        "note": "paths in, hashes out: reader threads pread() the tmpfs files into pinned slots, H2D, four stages"}
                    # the repo's C oracle (scalar port), one thread, on 256 MiB of shard 0
                    import xet_oracle
                    n = min(sizes[0], 256 << 20)
                    a = np.fromfile(paths[0], dtype=np.uint8, count=n)
                    t0 = time.perf_counter()
                    xet_oracle.hash_file(a)
                    cpu["port"] = {"value": round(n / (time.perf_counter() - t0) / 1e9, 4), "unit": "GB/s", "cores": 1,
                                   "sample": f"oracle/xet_oracle.c on the first {n >> 20} MiB of shard 0"}
        finally:
            shutil.rmtree(tmpdir, ignore_errors=True)
        all_ok = cx.all_true(ok)
        parity.update({"full_config": all_ok, "reference": kind,
                       "checked": (f"rank 0: all {len(sizes)} shards ({total / 1e9:.2f} GB) of the device-resident buffer that `value` times"
                                   + (f"; ranks 1..{world - 1}: their smallest shard ({min(sizes) / 1e9:.2f} GB) each" if world > 1 else "")),
                       "file_hashes_rank0": gpu_hex if rank == 0 else None})
        if not all_ok:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "value": None, "unit": "GB/s", "n_gpus": world, "parity": parity,
                                  "error": "file hashes differ from the CPU reference: no throughput reported"}))
            raise SystemExit(3)

    # ---- device-resident: `value`
    for _ in range(args.warmup):
        cl = step_device()
    ref_hashes = [f.file_hash for f in cl.files]
    cx.barrier()
    sampler = ClockSampler(cx.local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.5)  # let nvidia-smi come up so that its samples fall inside the timed region
    for _ in range(2):
        step_device()  # every rank (the step holds a collective): clocks back at their loaded level
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage = {k: 0.0 for k in ("gear_ms", "resolve_ms", "blake3_ms", "merkle_ms", "gear_kernel_ms", "leaf_kernel_ms")}
    launches = 0
    cx.barrier()
    ev0.record(cx.stream)
    for _ in range(args.steps):
        cl = step_device()
        launches += hasher.last_timings()["kernel_launches"]
        assert [f.file_hash for f in cl.files] == ref_hashes
    ev1.record(cx.stream)
    cx.barrier()
    ms = ev0.elapsed_time(ev1)
    # per-kernel durations for the roofline: the same step with the file groups run one after the other on one stream
    iso = ome_b200.XetHasher(device=cx.local_rank, stream=cx.stream.cuda_stream, stage_timing=True, overlap=False)
    iso_steps = 3
    for i in range(iso_steps + 1):
        cl_iso = iso.hash_buffers(buf, off_np, memory=ome_b200.hasher.MEM_DEVICE)
        if i:  # first call warms the workspace
            t = iso.last_timings()
            for k in stage:
                stage[k] += t[k]
    assert [f.file_hash for f in cl_iso.files] == ref_hashes
    iso.close()
    stage = {k: v / iso_steps for k, v in stage.items()}  # per step
    clocks = sampler.stop() if rank == 0 else None
    ms_max = cx.max_over_ranks(ms)
    n_chunks = sum(f.n_chunks for f in cl.files)
    value = total * world * args.steps / (ms_max / 1e3) / 1e9

    # ---- end to end: host (pinned) buffers through the same C-ABI call
    e2e, h2d_ceiling, host = None, None, None
    ok_pin = True
    if not args.no_e2e:
        try:
            host = torch.empty(total, dtype=torch.uint8, pin_memory=True)
            host.copy_(buf[:total])
            torch.cuda.synchronize()
        except RuntimeError as exc:  # not enough pinnable host memory for every rank's shard set
            print(f"[bench] rank {rank}: e2e skipped, cannot pin {total} bytes: {exc}", file=sys.stderr)
            host, ok_pin = None, False
    if not args.no_e2e and cx.all_true(ok_pin):
        def step_host():
            return hasher.hash_buffers(host, off_np, memory=ome_b200.hasher.MEM_HOST)

        cl_h = step_host()
        assert [f.file_hash for f in cl_h.files] == ref_hashes
        ems = cx.timed(step_host, args.steps, min(args.warmup, 3))
        e2e = {"value": round(total * world / (ems / 1e3) / 1e9, 3), "unit": "GB/s",
               "h2d_bytes_per_step": total * world, "d2h_bytes_per_step": (64 * len(sizes) + 12 * len(sizes) + 64) * world,
               "ms_per_step": round(ems, 3),
               "note": "pinned host shards -> xetgpu_hash_buffers(XETGPU_MEM_HOST): H2D copy, all four stages, "
                       "D2H of per-file hashes inside the timed region"}
        # the host-side ceiling of that path: the same pinned buffers copied by all ranks at once, no kernels
        def copy_only():
            buf[:total].copy_(host, non_blocking=True)

        cms = cx.timed(copy_only, 3, 1)
        h2d_ceiling = {"value": round(total * world / (cms / 1e3) / 1e9, 3), "unit": "GB/s",
                       "e2e_fraction_of_ceiling": round(e2e["value"] / (total * world / (cms / 1e3) / 1e9), 4),
                       "note": f"{world} rank(s) each copying their {total / 1e9:.2f} GB of pinned host memory to their GPU "
                               "at the same time, no kernels: what the host side of this box can feed"}
    del host

    record = None
    if rank == 0:
        peak, peak_src = peaks()
        leaf_ms, gear_ms = stage["leaf_kernel_ms"], stage["gear_kernel_ms"]
        achieved = total / (leaf_ms / 1e3) / 1e9 if leaf_ms > 0 else None
        ncu_path = NCU_SUMMARY if os.path.exists(NCU_SUMMARY) else NCU_SUMMARY_OLD
        traffic = traffic_src = pipes = int_issue = None
        if os.path.exists(ncu_path):
            ncu = json.load(open(ncu_path))
            rel = os.path.relpath(ncu_path, ROOT)
            ratio = ncu.get("leaf_group_kernel_traffic_per_algorithmic_byte")
            if ratio:
                traffic = round(ratio * total / 1e9, 3)
                traffic_src = f"{rel}: (dram read + write) / algorithmic bytes = {ratio}, scaled to this launch; GB"
            k3, k1 = ncu.get("kernels", {}).get("leaf_group_kernel", {}), ncu.get("kernels", {}).get("gear_scan_kernel", {})
            pipes = {"source": f"ncu --set full ({rel}), not measured in this run",
                     "alu_pipe_pct_of_peak": k3.get("alu_pct"), "fma_pipe_pct_of_peak": k3.get("fma_pct"),
                     "issue_slots_pct": k3.get("issue_pct"), "dram_pct_of_peak": k3.get("dram_pct")}
            # second roof: warp-instruction issue.  lane-instructions per byte are a property of the compiled kernels
            # (ncu smsp__inst_executed x 32 / algorithmic bytes); the machine issues 4 warp-instructions per clock per
            # SM = 128 lane-instructions (profiles/r1_int_pipes.txt: 64 ALU + 64 FMA lanes/clk/SM measured).
            alg = float(ncu.get("algorithmic_bytes_per_launch", 0)) or None
            try:
                ipb3 = float(str(k3.get("warp_insts")).split()[0]) * 32 / alg
                ipb1 = float(str(k1.get("warp_insts")).split()[0]) * 32 / alg
            except (TypeError, ValueError, ZeroDivisionError):
                ipb3 = ipb1 = None
            sm_hz = (clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0) * 1e6
            sms = torch.cuda.get_device_properties(cx.local_rank).multi_processor_count
            if ipb3 and achieved:
                issue_peak = 128.0 * sms * sm_hz  # lane-instructions per second
                pipeline_bps = total / (ms_max / args.steps / 1e3)
                int_issue = {
                    "bound": "int_issue", "unit": "lane-instr/s", "peak": issue_peak,
                    "peak_source": f"4 warp-instructions/clk/SM x 32 lanes x {sms} SMs x {sm_hz / 1e6:.0f} MHz (clock sampled in this run)",
                    "leaf_group_kernel": {"lane_instr_per_byte": round(ipb3, 3), "achieved": ipb3 * achieved * 1e9,
                                          "frac": round(ipb3 * achieved * 1e9 / issue_peak, 4)},
                    "gear_scan_kernel": {"lane_instr_per_byte": round(ipb1, 3),
                                         "frac": round(ipb1 * (total / (gear_ms / 1e3)) / issue_peak, 4) if gear_ms > 0 else None},
                    "pipeline": {"lane_instr_per_byte": round(ipb1 + ipb3, 3),
                                 "roof_gbs": round(issue_peak / (ipb1 + ipb3) / 1e9, 1),
                                 "frac": round((ipb1 + ipb3) * pipeline_bps / issue_peak, 4),
                                 "note": "both hot kernels' instructions against the issue rate: the ceiling for ANY schedule of "
                                         "these two instruction streams on one GPU, overlapped or not"}}
        roofline = {"bound": "hbm", "kernel": "leaf_group_kernel", "achieved": round(achieved, 1) if achieved else None,
                    "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4) if achieved else None,
                    "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": total, "binding_resource": pipes, "second_roof": int_issue,
                    "note": "the contract's roof is HBM (algorithmic bytes / kernel time / measured copy bandwidth); the resource "
                            "that actually binds is ALU-pipe issue (second_roof): ~15 lane-instructions per byte of keyed BLAKE3",
                    "per_stage_ms": {k: round(v, 4) for k, v in stage.items()},
                    "gear_kernel_achieved_gbs": round(total / (gear_ms / 1e3) / 1e9, 1) if gear_ms > 0 else None,
                    "pipeline_frac_of_hbm": round(value / world / peak, 4)}
        record = {
            "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_max / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "Llama-3-8B bf16 checkpoint shape: 4 safetensors-shaped shards per GPU, device-resident"
                                   + ("" if args.scale == 1.0 else f" (SCALED x{args.scale}: not the named config)"),
                       "shard_bytes": sizes, "bytes_per_gpu": total, "chunks_per_gpu": n_chunks,
                       "partition": "by file, one shard set per GPU" if world > 1 else "single GPU",
                       "cpu_affinity": f"each rank bound to the {cx.numa_cores} cores NVML reports local to its GPU" if cx.numa_cores else None,
                       "l2": "inputs (16 GB) larger than L2 (126 MB); no explicit flush"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "parity": parity, "h2d_ceiling": h2d_ceiling if world > 1 else None,
        }
        if world == 1 and h2d_ceiling:
            record["e2e"]["h2d_ceiling"] = h2d_ceiling
    hasher.close()
    del buf
    torch.cuda.empty_cache()
    return record


def leg_c1(cx: Ctx):
    """BASELINE config 1: one 1 GiB shard.  Contents = the repo's 1 GiB golden vector (16 x 64 MiB splitmix streams,
    seeds 100..115), whose hf_xet file hash is committed in tests/golden/xet_golden.json."""
    import numpy as np
    import torch
    import ome_b200
    name = "one_GiB_16x64MiB_splitmix_seeds_100_115"
    gold = next(c for c in json.load(open(os.path.join(ROOT, "tests", "golden", "xet_golden.json")))["cases"]
                if c["name"] == name)
    n = gold["n"]
    hasher = ome_b200.XetHasher(device=cx.local_rank, stream=cx.stream.cuda_stream, stage_timing=True)
    buf = torch.empty(n + 256, dtype=torch.uint8, device="cuda")
    pos = 0
    for seg in gold["spec"]:
        hasher.fill_splitmix(buf.data_ptr() + pos, seg[1], 0, seg[2])
        pos += seg[2]
    off = np.asarray([0, n], dtype=np.uint64)
    cl = hasher.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE)
    ok = cl.files[0].hex == gold["file_hash"] and cl.files[0].n_chunks == gold["n_chunks"]
    steps = max(5, min(cx.args.steps, 20))
    ms = cx.timed(lambda: hasher.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE), steps, 3)
    t = hasher.last_timings()
    hasher.close()
    return {"workload": "config 1: single 1 GiB shard, device-resident", "bytes": n, "chunks": int(cl.files[0].n_chunks),
            "value": round(n / ms / 1e6, 1) if ok else None, "unit": "GB/s", "ms_per_step": round(ms, 4), "steps": steps,
            "launches_per_step": t["kernel_launches"],
            "stage_ms": {k: round(t[k], 4) for k in ("gear_ms", "resolve_ms", "blake3_ms", "merkle_ms")},
            "parity": {"file_hash_equals_hf_xet_golden": ok, "golden": f"tests/golden/xet_golden.json: {name}"}}


def c5_sizes(n_files: int = 10000):
    import numpy as np
    import xet_inputs
    u = xet_inputs.splitmix_bytes(5, 8 * n_files).view("<u8").astype(np.float64) / 2.0**64
    s = np.exp(np.log(4096) + u * (np.log(64 << 20) - np.log(4096))).astype(np.int64)
    s[17] = 0       # one empty file
    s[18] = 100     # sub-8 KiB files: single chunk
    s[19] = 8191
    return [int(x) for x in s]


def leg_c5(cx: Ctx):
    """BASELINE config 5: 10,000 files, sizes log-uniform in [4 KiB, 64 MiB] (splitmix seed 5), contents seed
    50000 + i, packed in one device buffer; a subset is checked against the C oracle on the device bytes."""
    import numpy as np
    import torch
    import ome_b200
    import xet_oracle
    sizes = c5_sizes()
    if cx.args.scale != 1.0:
        sizes = sizes[: max(50, int(len(sizes) * cx.args.scale))]
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    total = int(off[-1])
    hasher = ome_b200.XetHasher(device=cx.local_rank, stream=cx.stream.cuda_stream, stage_timing=True)
    buf = torch.empty(total + 256, dtype=torch.uint8, device="cuda")
    for i, n in enumerate(sizes):
        if n:
            hasher.fill_splitmix(buf.data_ptr() + int(off[i]), 50000 + i, 0, n)
    torch.cuda.synchronize()
    cl = hasher.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE)
    # parity subset: every file below 64 KiB (incl. the empty and sub-8 KiB ones) and every 625th file whatever its size
    subset = sorted({i for i, n in enumerate(sizes) if n < 65536} | set(range(0, len(sizes), 625)))
    bad, sub_bytes = 0, 0
    for i in subset:
        a = buf[int(off[i]): int(off[i + 1])].cpu().numpy()
        sub_bytes += a.size
        bad += xet_oracle.hash_file(a).file_hash != cl.files[i].file_hash
    ok = bad == 0
    digest = hashlib.sha256(b"".join(f.file_hash for f in cl.files)).hexdigest()
    steps = 5
    ms = cx.timed(lambda: hasher.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE), steps, 2)
    t = hasher.last_timings()
    hasher.close()
    return {"workload": f"config 5: {len(sizes)} files, log-uniform 4 KiB-64 MiB, one empty and three sub-8 KiB files, device-resident",
            "scale_factor": 1.0 if cx.args.scale == 1.0 else cx.args.scale,
            "files": len(sizes), "bytes": total, "chunks": int(cl.files.records["n_chunks"].sum()),
            "value": round(total / ms / 1e6, 1) if ok else None, "unit": "GB/s", "ms_per_step": round(ms, 3), "steps": steps,
            "launches_per_step": t["kernel_launches"], "launches_per_file": round(t["kernel_launches"] / len(sizes), 4),
            "parity": {"oracle_equal": ok, "files_checked": len(subset), "bytes_checked": sub_bytes,
                       "subset": "all files < 64 KiB + every 625th file, device bytes -> oracle/xet_oracle.c",
                       "sha256_of_all_file_hashes": digest}}


def leg_config3(cx: Ctx):
    """BASELINE config 3: 30 shards (~138 GB) file-partitioned over the ranks, LPT (strong scaling)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import ome_b200
    from ome_b200 import sharded
    sizes, _ = shard_layout(C3_SIZES, cx.args.scale)
    part = sharded.FilePartition.lpt(sizes, cx.world)
    mine = part.assignment[cx.rank]
    hasher = ome_b200.XetHasher(device=cx.local_rank, stream=cx.stream.cuda_stream)
    offs = np.concatenate([[0], np.cumsum([sizes[i] for i in mine])]).astype(np.uint64)
    buf = torch.empty(int(offs[-1]) + 256, dtype=torch.uint8, device="cuda")
    for k, i in enumerate(mine):
        fill_file_range(hasher, buf, int(offs[k]), sizes[i], C3_SEED + i)
    torch.cuda.synchronize()
    hashes = torch.zeros(len(sizes), 32, dtype=torch.int32, device="cuda")

    def step():
        hashes.zero_()
        if mine:
            cl = hasher.hash_buffers(buf, offs, memory=ome_b200.hasher.MEM_DEVICE)  # this rank's shards in one batched call
            mineh = np.frombuffer(b"".join(f.file_hash for f in cl.files), dtype=np.uint8).reshape(-1, 32).astype(np.int32)
            hashes[torch.as_tensor(mine, device="cuda")] = torch.from_numpy(mineh).to("cuda")
        dist.all_reduce(hashes)  # disjoint rows: a sum is a gather; every rank ends with all 30 file hashes

    step()
    all_hashes = hashes.to(torch.uint8).cpu().numpy()
    # parity: every rank hands the first shard it owns (device bytes -> tmpfs) to hf_xet; the owner of the short last
    # shard checks that one too
    which = mine[:1] + ([len(sizes) - 1] if len(sizes) - 1 in mine[1:] else [])
    ok, kind = True, "reference"
    if which and not cx.args.no_cpu and tmp_free_bytes() > cx.world * 2 * max(sizes):
        tmpdir = tempfile.mkdtemp(prefix=f"xetbench-c3-r{cx.rank}-", dir=tmp_root())
        try:
            paths = []
            for i in which:
                p = os.path.join(tmpdir, f"c3-{i}.safetensors")
                device_file_to_tmpfs(buf, int(offs[mine.index(i)]), sizes[i], p)
                paths.append(p)
            _, ref_hex, kind = cpu_hash_files(paths)
            ok = ref_hex == [ome_b200.hash_to_hex(all_hashes[i].tobytes()) for i in which]
        finally:
            shutil.rmtree(tmpdir, ignore_errors=True)
    else:
        which = []
    checked = [None] * cx.world
    dist.all_gather_object(checked, which)
    all_ok = cx.all_true(ok)
    steps = max(5, min(cx.args.steps, 10))
    ms = cx.timed(step, steps, 2)
    hasher.close()
    del buf
    torch.cuda.empty_cache()
    total = sum(sizes)
    if cx.rank != 0:
        return None
    return {"workload": "config 3: Llama-3-70B bf16 checkpoint shape, 30 safetensors-shaped shards, file-partitioned (LPT), device-resident"
                        + ("" if cx.args.scale == 1.0 else f" (SCALED x{cx.args.scale})"),
            "scaling": "strong", "files": len(sizes), "bytes": total, "value": round(total / ms / 1e6, 1) if all_ok else None,
            "unit": "GB/s", "ms_per_step": round(ms, 3), "steps": steps, "timing": "CUDA events on each rank's stream, barrier both sides, max over ranks, mean of the steps",
            "files_per_rank": [len(a) for a in part.assignment], "max_rank_bytes": max(part.loads),
            "balance": round(total / cx.world / max(part.loads), 4),
            "parity": {"equal": all_ok, "reference": kind, "shards_checked": sorted(i for w in checked for i in w),
                       "note": "each rank's first shard (and the short last shard): device bytes -> tmpfs -> hf_xet.hash_files, full size",
                       "sha256_of_all_file_hashes": hashlib.sha256(all_hashes.tobytes()).hexdigest()}}


def leg_config4(cx: Ctx):
    """BASELINE config 4: 163 shards (~689 GB) laid end to end and cut into one byte range per GPU; a shard that
    straddles a range boundary is shared by the two neighbouring ranks (halo + carried cut over NCCL P2P, fix-up
    window, merkle combine on the lower rank)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import ome_b200
    from ome_b200 import sharded
    sizes, _ = shard_layout(C4_SIZES, cx.args.scale)
    if cx.args.scale != 1.0:
        sizes = sizes[: max(2 * cx.world + 1, int(len(sizes) * min(1.0, cx.args.scale * 8))) | 1]  # odd: a shard straddles the middle
    n_files = len(sizes)
    plan = sharded.plan_concatenated(sizes, cx.world)
    owners, shared = sharded.shared_files(plan)
    groups = {f: dist.new_group(ranks=owners[f]) for f in shared}  # every rank creates every group, same order
    lay = sharded.layout_pieces(plan[cx.rank])
    hasher = ome_b200.XetHasher(device=cx.local_rank, stream=cx.stream.cuda_stream)
    buf = torch.empty(lay.nbytes, dtype=torch.uint8, device="cuda")
    for k, p in enumerate(lay.pieces):
        if p.end > p.begin:
            fill_file_range(hasher, buf, lay.data_offset(k), p.file_len, C4_SEED + p.file, p.begin, p.end)
    torch.cuda.synchronize()
    state = {"hashes": None, "rehashed": 0}
    phases = {}

    def step(ph=None):
        h, r = sharded.hash_concatenated(hasher, buf, lay, plan, n_files, groups, comm_device=cx.dev, phases=ph,
                                         memory=ome_b200.hasher.MEM_DEVICE)
        state["hashes"], state["rehashed"] = h, max(state["rehashed"], r)

    step()
    all_hashes = state["hashes"]
    # parity: every straddling shard is reassembled in tmpfs from the DEVICE bytes of its owners and hashed whole by
    # hf_xet on the lower owner; rank 0 also checks the first shard it holds whole.
    ok, kind, checked = True, "reference", []
    do_parity = not cx.args.no_cpu and tmp_free_bytes() > (len(shared) + 2) * max(sizes) * 1.2
    tmpdir = cx.shared_tmpdir("c4") if do_parity else None
    if do_parity:
        try:
            for k, p in enumerate(lay.pieces):
                if p.file in shared and p.end > p.begin:
                    device_file_to_tmpfs(buf, lay.data_offset(k), p.end - p.begin, os.path.join(tmpdir, f"c4-{p.file}.bin"), p.begin)
            cx.barrier()
            mine = [f for f in shared if owners[f][0] == cx.rank]
            if cx.rank == 0:
                k0 = next((k for k, p in enumerate(lay.pieces) if p.whole and p.file_len), None)
                if k0 is not None:
                    f0 = lay.pieces[k0].file
                    device_file_to_tmpfs(buf, lay.data_offset(k0), sizes[f0], os.path.join(tmpdir, f"c4-{f0}.bin"))
                    mine = mine + [f0]
            if mine:
                _, ref_hex, kind = cpu_hash_files([os.path.join(tmpdir, f"c4-{f}.bin") for f in mine])
                ok = ref_hex == [ome_b200.hash_to_hex(all_hashes[f].tobytes()) for f in mine]
            checked = mine
            cx.barrier()
        finally:
            if cx.rank == 0:
                shutil.rmtree(tmpdir, ignore_errors=True)
    gathered = [None] * cx.world
    dist.all_gather_object(gathered, checked)
    all_ok = cx.all_true(ok)
    steps = max(5, min(cx.args.steps, 8))
    ms = cx.timed(lambda: step(phases), steps, 1)
    # phases hold host wall-clock sums over warm-up + timed steps on this rank; report the slowest rank's per-step means
    # (a fixed key list keeps the reductions identical on every rank: a rank without a shared shard has no such phases)
    keys = ("halo_exchange_ms", "batched_pass_ms", "carry_chain_ms", "gather_ms", "merkle_ms", "result_allreduce_ms")
    ph = {k: cx.max_over_ranks(phases.get(k, 0.0) / (steps + 1)) for k in keys}
    rehashed = int(cx.max_over_ranks(float(state["rehashed"])))
    hasher.close()
    del buf
    torch.cuda.empty_cache()
    total = sum(sizes)
    if cx.rank != 0:
        return None
    return {"workload": "config 4: DeepSeek-V3 fp8 checkpoint shape, 163 safetensors-shaped shards laid end to end, byte-range partitioned, device-resident"
                        + ("" if cx.args.scale == 1.0 else f" (SCALED x{cx.args.scale}, {n_files} shards)"),
            "scaling": "strong", "files": n_files, "bytes": total, "bytes_per_rank": [sum(p.end - p.begin for p in r) for r in plan],
            "value": round(total / ms / 1e6, 1) if all_ok else None, "unit": "GB/s", "ms_per_step": round(ms, 3), "steps": steps,
            "timing": "CUDA events on each rank's stream, barrier both sides, max over ranks, mean of the steps",
            "shared_files": shared, "max_rehashed_bytes": rehashed,
            "phase_ms_slowest_rank": {k: round(v, 3) for k, v in ph.items()},
            "parity": {"equal": all_ok, "reference": kind, "shards_checked": sorted(f for w in gathered for f in w),
                       "note": "all straddling shards (reassembled in tmpfs from the device bytes of both owners) and rank 0's first whole "
                               "shard: hf_xet.hash_files at full size" if do_parity else "skipped: not enough tmpfs",
                       "sha256_of_all_file_hashes": hashlib.sha256(all_hashes.tobytes()).hexdigest()}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=float, default=float(os.environ.get("XETBENCH_SCALE", "1.0")),
                    help="shrink every shard by this factor (smoke runs only; 1.0 = the named config)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip every CPU leg (parity gates included)")
    ap.add_argument("--only", default="", help="comma list of legs to run besides the headline: c1,c5,c3,c4 (default: all that apply)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    cx = Ctx(args)
    legs = set(args.only.split(",")) if args.only else {"c1", "c5", "c3", "c4"}
    record = leg_config2(cx)
    extra = {}
    try:
        if cx.world == 1:
            configs = {}
            if "c1" in legs:
                configs["c1"] = leg_c1(cx)
            if "c5" in legs:
                configs["c5"] = leg_c5(cx)
            extra["configs"] = configs
        else:
            if "c3" in legs:
                extra["config3"] = leg_config3(cx)
            if "c4" in legs and (cx.world == 8 or args.scale != 1.0):
                extra["config4"] = leg_config4(cx)
    except Exception as exc:  # the headline number stands on its own; a failed side leg is reported, not hidden
        import traceback
        traceback.print_exc()
        extra["side_leg_error"] = f"rank {cx.rank}: {type(exc).__name__}: {exc}"
    if cx.rank == 0:
        record.update(extra)
        print(json.dumps(record))
    if cx.world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
