#!/usr/bin/env python
"""bench.py — GB/s of model bytes chunked + chunk-hashed + file-hashed, bit-exact, on N B200s.

    python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): the Llama-3-8B bf16 checkpoint shape — four safetensors-shaped
shards of 4.98 / 5.00 / 4.92 / 1.17 GB (16.06 GB) of counter-based splitmix64 bytes.
At N > 1 every rank holds its own copy of that shape with different seeds (file-partitioned, weak
scaling: no data-path collective; only the 32-byte file hashes are gathered to rank 0).

One step = one pass of the whole hot path (K1 gear candidates -> K2 boundary resolution -> K3 keyed
BLAKE3 chunk hashes -> K4 merkle + file hash) over all shards of the rank.
  value : inputs already resident in HBM, timed with CUDA events on the stream the library launches on
  e2e   : the same call through the C ABI with the shards in pinned HOST memory (H2D inside the timed region)
Inputs (16 GB per rank) are far larger than the 126 MB L2, so no explicit L2 flush is needed.

`--impl reference` times the reference's own CPU arithmetic: hf_xet.hash_files — the official compiled
build of huggingface/xet-core (the crate /root/reference/pkg/xet pins in Cargo.lock; its Rust sources are
not vendored and cannot be compiled here) — on a bounded sample of the same shards in tmpfs.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "GB/s model bytes chunked+hashed bit-exact"
# real Llama-3-8B safetensors shard sizes (model-0000{1..4}-of-00004.safetensors)
LLAMA3_8B_SHARDS = [4_976_698_672, 4_999_802_720, 4_915_916_176, 1_168_138_808]
SEED_BASE = 10


def safetensors_header(payload_bytes: int) -> bytes:
    """8-byte LE header length + JSON header padded to 8 bytes (pkg/hfutil/modelconfig/safetensors.go:78-100)."""
    n_elems = payload_bytes // 2
    hdr = json.dumps({"__metadata__": {"format": "pt"},
                      "t0": {"dtype": "BF16", "shape": [n_elems], "data_offsets": [0, n_elems * 2]}},
                     separators=(",", ":")).encode()
    hdr += b" " * (-len(hdr) % 8)
    return len(hdr).to_bytes(8, "little") + hdr


def shard_layout(scale: float):
    sizes = [max(1 << 20, int(s * scale)) for s in LLAMA3_8B_SHARDS]
    offsets = [0]
    for s in sizes:
        offsets.append(offsets[-1] + s)
    return sizes, offsets


def fill_shards(hasher, buf, sizes, offsets, seed_base):
    """Write the safetensors-shaped shards into the device buffer (header via H2D, payload on device)."""
    import torch
    for i, size in enumerate(sizes):
        hdr = safetensors_header(size - 256)
        hdr_t = torch.frombuffer(bytearray(hdr), dtype=torch.uint8)
        buf[offsets[i]:offsets[i] + len(hdr)].copy_(hdr_t)
        hasher.fill_splitmix(buf.data_ptr() + offsets[i] + len(hdr), seed_base + i, 0, size - len(hdr))
    torch.cuda.synchronize()


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


# ------------------------------------------------------------------------------------------- CPU side
def cpu_sample_files(host_views, sizes, sample_cap_bytes, tmpdir):
    """Write a bounded prefix of every shard to tmpfs; returns (paths, bytes)."""
    paths, total = [], 0
    per = sample_cap_bytes // len(sizes)
    for i, (view, size) in enumerate(zip(host_views, sizes)):
        n = min(size, per)
        p = os.path.join(tmpdir, f"shard-{i}.safetensors")
        view[:n].numpy().tofile(p)
        paths.append(p)
        total += n
    return paths, total


def time_hf_xet(paths):
    import hf_xet
    t0 = time.perf_counter()
    infos = hf_xet.hash_files(paths)
    return time.perf_counter() - t0, [i.hash for i in infos]


def tmp_root():
    for d in ("/dev/shm", tempfile.gettempdir()):
        if os.path.isdir(d) and os.access(d, os.W_OK):
            return d
    return "."


def host_shards_numpy(sizes, seed_base, cap_each):
    """Host-generated shard prefixes (numpy splitmix64) for the reference arm — no GPU involved."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import xet_inputs
    out = []
    for i, size in enumerate(sizes):
        hdr = np.frombuffer(safetensors_header(size - 256), dtype=np.uint8)
        n = min(size, cap_each)
        body = xet_inputs.splitmix_bytes(seed_base + i, max(0, n - hdr.size))
        out.append(np.concatenate([hdr, body])[:n])
    return out


def run_reference(args):
    """Reference arm: hf_xet.hash_files (compiled xet-core) on the host cores; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sizes, _ = shard_layout(args.scale)
    cap_each = int(args.cpu_sample_gb * 1e9) // len(sizes)
    shards = host_shards_numpy(sizes, SEED_BASE, cap_each)
    tmpdir = tempfile.mkdtemp(prefix="xetbench-", dir=tmp_root())
    try:
        paths, total = [], 0
        for i, a in enumerate(shards):
            p = os.path.join(tmpdir, f"shard-{i}.safetensors")
            a.tofile(p)
            paths.append(p)
            total += a.size
        for _ in range(args.warmup):
            time_hf_xet(paths)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            time_hf_xet(paths)
        dt = time.perf_counter() - t0
    finally:
        shutil.rmtree(tmpdir, ignore_errors=True)
    value = total * args.steps / dt / 1e9
    sample = (f"hf_xet 1.5.0 hash_files (compiled huggingface/xet-core) on the first {cap_each / 1e9:.2f} GB of each of the "
              f"{len(sizes)} Llama-3-8B-shaped shards ({total / 1e9:.2f} GB) in tmpfs, one call per step")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(value, 4), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "Llama-3-8B bf16 checkpoint shape, 4 safetensors-shaped shards (bounded sample)",
                   "shard_bytes": sizes, "sample_bytes": total},
        "cpu_baseline": {"value": round(value, 4), "unit": "GB/s", "cores": os.cpu_count(), "kind": "reference",
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
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=float, default=float(os.environ.get("XETBENCH_SCALE", "1.0")),
                    help="shrink every shard by this factor (smoke runs only; 1.0 = the named config)")
    ap.add_argument("--cpu-sample-gb", type=float, default=6.0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import ome_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    numa_cores = bind_to_gpu_numa(local_rank) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sizes, offsets = shard_layout(args.scale)
    total = offsets[-1]
    stream = torch.cuda.Stream()  # the library launches on this stream; the timing events are recorded on it
    torch.cuda.set_stream(stream)
    hasher = ome_b200.XetHasher(device=local_rank, stream=stream.cuda_stream, stage_timing=True)
    buf = torch.empty(total + 256, dtype=torch.uint8, device="cuda")
    fill_shards(hasher, buf, sizes, offsets, SEED_BASE + 100 * rank)
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

    # ---- device-resident: `value`
    for _ in range(args.warmup):
        cl = step_device()
    ref_hashes = [f.file_hash for f in cl.files]
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.5)  # let nvidia-smi come up so that its samples fall inside the timed region
    for _ in range(2):
        step_device()  # every rank (the step holds a collective): clocks back at their loaded level
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage = {k: 0.0 for k in ("gear_ms", "resolve_ms", "blake3_ms", "merkle_ms", "gear_kernel_ms", "leaf_kernel_ms")}
    launches = 0
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        cl = step_device()
        launches += hasher.last_timings()["kernel_launches"]
        assert [f.file_hash for f in cl.files] == ref_hashes
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    # per-kernel durations for the roofline: the same step with the file groups run one after the other on one
    # stream (in the timed loop above two groups overlap on two streams, which stretches each kernel's own time)
    iso = ome_b200.XetHasher(device=local_rank, stream=stream.cuda_stream, stage_timing=True, overlap=False)
    iso_steps = 3
    for i in range(iso_steps + 1):
        cl_iso = iso.hash_buffers(buf, off_np, memory=ome_b200.hasher.MEM_DEVICE)
        if i:  # first call warms the workspace
            t = iso.last_timings()
            for k in stage:
                stage[k] += t[k]
    assert [f.file_hash for f in cl_iso.files] == ref_hashes
    iso.close()
    stage = {k: v * args.steps / iso_steps for k, v in stage.items()}  # reported per step below
    clocks = sampler.stop() if rank == 0 else None
    tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    n_chunks = sum(f.n_chunks for f in cl.files)
    value = total * world * args.steps / (ms_max / 1e3) / 1e9

    # ---- end to end: host (pinned) buffers through the same C-ABI call
    e2e = None
    host = None
    e2e_ok = torch.ones(1, dtype=torch.int32, device="cuda")
    if not args.no_e2e:
        try:
            host = torch.empty(total, dtype=torch.uint8, pin_memory=True)
            host.copy_(buf[:total])
            torch.cuda.synchronize()
        except RuntimeError as exc:  # not enough pinnable host memory for every rank's shard set
            print(f"[bench] rank {rank}: e2e skipped, cannot pin {total} bytes: {exc}", file=sys.stderr)
            host = None
            e2e_ok.zero_()
    if world > 1:
        dist.all_reduce(e2e_ok, op=dist.ReduceOp.MIN)   # all ranks or none
    if not args.no_e2e and int(e2e_ok.item()) == 1:

        def step_host():
            return hasher.hash_buffers(host, off_np, memory=ome_b200.hasher.MEM_HOST)

        for _ in range(min(args.warmup, 3)):
            cl_h = step_host()
        assert [f.file_hash for f in cl_h.files] == ref_hashes
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            cl_h = step_host()
        e1.record(stream)
        barrier()
        ems = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ems, op=dist.ReduceOp.MAX)
        e2e = {"value": round(total * world * args.steps / (float(ems.item()) / 1e3) / 1e9, 3), "unit": "GB/s",
               "h2d_bytes_per_step": total * world, "d2h_bytes_per_step": (64 * len(sizes) + 12 * len(sizes) + 64) * world,
               "ms_per_step": round(float(ems.item()) / args.steps, 3),
               "note": "pinned host shards -> xetgpu_hash_buffers(XETGPU_MEM_HOST): H2D copy, all four stages, "
                       "D2H of per-file hashes inside the timed region"}

    # ---- CPU baseline (rank 0, N=1 only): the reference arithmetic on the host cores, bounded sample
    cpu = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu:
        tmpdir = tempfile.mkdtemp(prefix="xetbench-", dir=tmp_root())
        try:
            if host is None:
                host = torch.empty(total, dtype=torch.uint8)
                host.copy_(buf[:total])
            views = [host[offsets[i]:offsets[i + 1]] for i in range(len(sizes))]
            paths, sample_bytes = cpu_sample_files(views, sizes, int(args.cpu_sample_gb * 1e9), tmpdir)
            time_hf_xet(paths[-1:])  # warm the library
            dt, ref_hex = time_hf_xet(paths)
            # parity on the very bytes the CPU hashed: the GPU path through hash_files on the same files
            gpu = hasher.hash_files(paths)
            t0 = time.perf_counter()
            gpu = hasher.hash_files(paths)
            dt_gpu = time.perf_counter() - t0
            parity = {"files": len(paths), "bytes": sample_bytes,
                      "file_hashes_equal_hf_xet": [f.hex for f in gpu.files] == ref_hex,
                      "same_files_through_xetgpu_hash_files": {
                          "value": round(sample_bytes / dt_gpu / 1e9, 2), "unit": "GB/s",
                          "note": "paths in, hashes out: reader threads pread() the tmpfs files into pinned slots, "
                                  "H2D, four stages; the same call the reference arm makes on the same files"}}
            cpu = {"value": round(sample_bytes / dt / 1e9, 4), "unit": "GB/s", "cores": os.cpu_count(),
                   "kind": "reference",
                   "sample": f"hf_xet 1.5.0 hash_files (compiled huggingface/xet-core; one call, file-parallel) on the first "
                             f"{sample_bytes / len(paths) / 1e9:.2f} GB of each of the {len(paths)} shards "
                             f"({sample_bytes / 1e9:.2f} GB, tmpfs), {dt:.2f} s wall"}
            # the repo's C oracle (scalar port), one thread, on 256 MiB of shard 0
            import xet_oracle
            n = min(sizes[0], 256 << 20)
            a = views[0][:n].numpy()
            t0 = time.perf_counter()
            xet_oracle.hash_file(a)
            cpu["port"] = {"value": round(n / (time.perf_counter() - t0) / 1e9, 4), "unit": "GB/s", "cores": 1,
                           "sample": f"oracle/xet_oracle.c on the first {n >> 20} MiB of shard 0"}
        finally:
            shutil.rmtree(tmpdir, ignore_errors=True)

    if rank == 0:
        peak, peak_src = peaks()
        # dominant kernel: leaf_group_kernel (keyed BLAKE3 leaves).  Algorithmic bytes per launch = every
        # input byte read once = the rank's total bytes (DESIGN.md "Roofline").
        leaf_ms = stage["leaf_kernel_ms"] / args.steps
        achieved = total / (leaf_ms / 1e3) / 1e9 if leaf_ms > 0 else None
        traffic, traffic_src, pipes = None, None, None
        ncu_path = os.path.join(ROOT, "profiles", "r1_ncu_summary.json")
        if os.path.exists(ncu_path):  # DRAM bytes per algorithmic byte of this kernel, from the committed ncu --set full capture
            ncu = json.load(open(ncu_path))
            ratio = ncu.get("leaf_group_kernel_traffic_per_algorithmic_byte")
            if ratio:
                traffic = round(ratio * total / 1e9, 3)
                traffic_src = f"profiles/r1_ncu_summary.json: (dram read + write) / algorithmic bytes = {ratio} on a 2 GiB launch, scaled to this launch; GB"
            k = ncu.get("kernels", {}).get("leaf_group_kernel", {})
            pipes = {"source": "ncu --set full (profiles/r1_ncu_summary.json), not measured in this run",
                     "alu_pipe_pct_of_peak": k.get("alu_pct"), "fma_pipe_pct_of_peak": k.get("fma_pct"),
                     "issue_slots_pct": k.get("issue_pct"), "dram_pct_of_peak": k.get("dram_pct")}
        roofline = {"bound": "hbm", "kernel": "leaf_group_kernel", "achieved": round(achieved, 1) if achieved else None,
                    "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4) if achieved else None,
                    "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": total, "binding_resource": pipes,
                    "note": "ALU-issue bound, not HBM bound: ncu shows the ALU pipe 89 % busy at 24 % DRAM throughput "
                            "(profiles/r1_summary.md); ~15 lane-instructions per byte of keyed BLAKE3",
                    "per_stage_ms": {k: round(v / args.steps, 4) for k, v in stage.items()},
                    "gear_kernel_achieved_gbs": round(total / (stage["gear_kernel_ms"] / args.steps / 1e3) / 1e9, 1)
                    if stage["gear_kernel_ms"] > 0 else None}
        print(json.dumps({
            "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_max / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "Llama-3-8B bf16 checkpoint shape: 4 safetensors-shaped shards per GPU, device-resident"
                                   + ("" if args.scale == 1.0 else f" (SCALED x{args.scale}: not the named config)"),
                       "shard_bytes": sizes, "bytes_per_gpu": total, "chunks_per_gpu": n_chunks,
                       "partition": "by file, one shard set per GPU" if world > 1 else "single GPU",
                       "cpu_affinity": f"each rank bound to the {numa_cores} cores NVML reports local to its GPU" if numa_cores else None,
                       "l2": "inputs (16 GB) larger than L2 (126 MB); no explicit flush"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "parity": parity,
        }))
    hasher.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
