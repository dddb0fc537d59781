"""Multi-GPU host logic: one process per GPU, `torch.distributed` for the (tiny) exchanges.

The path shards in two ways:

* by file  — every rank hashes whole files; nothing is exchanged but the per-file results
  (BASELINE config 3).  `FilePartition` + `hash_files_partitioned`.
* by byte range within one file — rank r owns bytes [begin_r, end_r) (BASELINE config 4).
  Chunk boundaries are a carry chain across ranges, so `hash_file_by_ranges` speculates:
    1. neighbours exchange the `RANGE_LEAD` bytes in front of each range (the 64-byte gear halo
       plus one maximum chunk, so that the chunk straddling the range start can be hashed by
       the rank that owns its end);
    2. every rank hashes its whole range at once, pretending a cut sits exactly at its range
       begin (rank 0 needs no pretence);
    3. the true carry (8 bytes: the last cut at or before the range begin) then travels down the
       ranks; a rank re-hashes only a small window at the front of its range from the true
       carry and splices it onto its speculative chunk list at the first common cut — from a
       shared cut on, the two chains are identical.  If the window holds no common cut it is
       widened, in the worst case to the whole range (still exact);
    4. rank 0 receives every rank's (hash, size) list and computes the merkle root / file hash.
  The bulk work (2) runs on all ranks in parallel; the chain (3) costs a few small launches per hop.

The engine is whatever object offers `hash_range`, `merkle_from_chunks` and `hash_buffers` with
the signatures of `ome_b200.XetHasher` — the product passes a XetHasher (CUDA, no CPU fallback);
the CPU tests drive the same code over gloo with an oracle-backed engine of their own.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

MAX_CHUNK = 131072
RANGE_LEAD = MAX_CHUNK + 64
FIX_WINDOW = 4 << 20   # first re-hash window after the true carry arrives


# --------------------------------------------------------------------------------- partitioning
@dataclass
class FilePartition:
    assignment: List[List[int]]     # per rank: file indices, in processing order
    loads: List[int]                # per rank: bytes

    @staticmethod
    def lpt(sizes: Sequence[int], world: int) -> "FilePartition":
        """Longest-processing-time-first: biggest file to the currently lightest rank."""
        assignment = [[] for _ in range(world)]
        loads = [0] * world
        for i in sorted(range(len(sizes)), key=lambda i: (-sizes[i], i)):
            r = min(range(world), key=lambda r: (loads[r], r))
            assignment[r].append(i)
            loads[r] += sizes[i]
        return FilePartition(assignment, loads)


def split_byte_ranges(file_len: int, world: int, align: int = 4096) -> List[Tuple[int, int]]:
    """`world` contiguous ranges covering [0, file_len); interior bounds aligned down to `align`.
    Every range except possibly the last is at least RANGE_LEAD long when the file allows it;
    trailing ranks get empty ranges for very small files."""
    if world <= 0:
        raise ValueError("world must be positive")
    per = max(-(-file_len // world), RANGE_LEAD + align)
    bounds = [0]
    for r in range(1, world):
        b = min(file_len, (r * per) // align * align)
        if file_len - b < RANGE_LEAD:      # do not leave a sliver for the last ranks: fold it into this one
            b = file_len
        bounds.append(max(b, bounds[-1]))
    bounds.append(file_len)
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


@dataclass
class Piece:
    file: int      # global file index
    begin: int     # byte range of the file this rank owns
    end: int
    file_len: int

    @property
    def whole(self) -> bool:
        return self.begin == 0 and self.end == self.file_len


def plan_concatenated(file_sizes: Sequence[int], world: int, align: int = 4096) -> List[List[Piece]]:
    """BASELINE config 4: the files laid end to end are cut into `world` equal byte ranges; a file that
    straddles a range boundary is shared by consecutive ranks (byte-range protocol), every other file is
    hashed whole by one rank.  A boundary that would leave a piece shorter than RANGE_LEAD + align at either
    end of a file snaps to that end, and in-file boundaries are aligned down to `align`.
    Returns, per rank, its pieces in file order; every file's pieces are consecutive ranks' and tile it."""
    total = sum(file_sizes)
    starts = np.concatenate([[0], np.cumsum(file_sizes)]).astype(np.int64)
    min_piece = RANGE_LEAD + align
    cuts = [0]                                    # positions in the concatenated space
    for r in range(1, world):
        pos = total * r // world
        f = int(np.searchsorted(starts, pos, side="right") - 1)
        f = min(max(f, 0), len(file_sizes) - 1)
        off = (pos - int(starts[f])) // align * align
        if off < min_piece:
            off = 0
        elif file_sizes[f] - off < min_piece:
            off = file_sizes[f]
        cuts.append(max(int(starts[f]) + off, cuts[-1]))
    cuts.append(total)
    plan: List[List[Piece]] = [[] for _ in range(world)]
    for r in range(world):
        lo, hi = cuts[r], cuts[r + 1]
        for f, n in enumerate(file_sizes):
            fb, fe = int(starts[f]), int(starts[f]) + n
            if n == 0:
                if lo <= fb < hi or (r == world - 1 and fb == total):   # an empty file belongs to one rank
                    plan[r].append(Piece(f, 0, 0, 0))
                continue
            b, e = max(lo, fb), min(hi, fe)
            if b < e:
                plan[r].append(Piece(f, b - fb, e - fb, n))
    return plan


# ------------------------------------------------------------------------------- file partition
def hash_files_partitioned(engine, local_files: Sequence[Tuple[int, object]], n_files: int,
                           group=None, comm_device: Optional[torch.device] = None):
    """Every rank hashes its own files [(global index, bytes-like)], then all ranks receive all
    results.  Returns (file_hashes uint8 [n_files, 32], n_chunks int64 [n_files])."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    dev = comm_device or torch.device("cpu")
    hashes = torch.zeros(n_files, 32, dtype=torch.uint8)
    counts = torch.zeros(n_files, dtype=torch.int64)
    for idx, data in local_files:
        n = _nbytes(data)
        cl = engine.hash_buffers(data, [0, n])
        hashes[idx] = torch.frombuffer(bytearray(cl.files[0].file_hash), dtype=torch.uint8)
        counts[idx] = cl.files[0].n_chunks
    if world > 1:
        # disjoint rows are non-zero on exactly one rank: a sum all-reduce is a gather
        h = hashes.to(torch.int32).to(dev)
        c = counts.to(dev)
        dist.all_reduce(h, group=group)
        dist.all_reduce(c, group=group)
        hashes, counts = h.to("cpu").to(torch.uint8), c.to("cpu")
    return hashes.numpy(), counts.numpy()


def _nbytes(data) -> int:
    if hasattr(data, "numel"):
        return data.numel() * data.element_size()
    return np.asarray(data).nbytes if not isinstance(data, (bytes, bytearray, memoryview)) else len(data)


# --------------------------------------------------------------------------------- byte ranges
@dataclass
class RangeOutcome:
    chunk_ends: np.ndarray       # file-relative exclusive ends of the chunks this rank owns
    chunk_hashes: np.ndarray     # [n, 32]
    exit_cut: int
    rehashed_bytes: int          # bytes of fix-up window(s) hashed after the true carry arrived
    merkle_root: Optional[bytes] = None   # rank 0 only
    file_hash: Optional[bytes] = None     # rank 0 only
    n_chunks_total: Optional[int] = None  # rank 0 only


def exchange_lead(piece: torch.Tensor, lead_capacity: int, ranges: Sequence[Tuple[int, int]], group=None) -> int:
    """Halo exchange.  `piece` = [lead_capacity bytes of headroom][this rank's own range bytes]; the
    call fills the last `lead` bytes of the headroom with the bytes that precede the range, received
    from the previous rank, and sends this rank's last RANGE_LEAD bytes to the next one.  Ranges come
    from split_byte_ranges: every non-empty range that has a non-empty successor is at least
    RANGE_LEAD long, so one neighbour always holds the whole halo.  Returns lead."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    begin, end = ranges[rank]
    n_own = end - begin
    lead = min(begin, RANGE_LEAD) if n_own else 0
    if lead > lead_capacity:
        raise ValueError("headroom in front of the range is smaller than the halo")
    # gloo moves host memory only: stage device tensors through the host when the group runs on it (single-GPU tests)
    host_hop = piece.is_cuda and dist.get_backend(group) == "gloo"
    ops = []
    nxt = ranges[rank + 1] if rank + 1 < world else (0, 0)
    if nxt[1] > nxt[0]:                                    # the next rank needs my tail
        if n_own < RANGE_LEAD:
            raise ValueError("a range followed by a non-empty range must be at least RANGE_LEAD bytes")
        tail = piece[lead_capacity + n_own - RANGE_LEAD: lead_capacity + n_own].contiguous()
        ops.append(dist.P2POp(dist.isend, tail.cpu() if host_hop else tail, _peer(rank + 1, group), group))
    buf = None
    if lead:
        buf = torch.empty(RANGE_LEAD, dtype=torch.uint8, device="cpu" if host_hop else piece.device)
        ops.append(dist.P2POp(dist.irecv, buf, _peer(rank - 1, group), group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    if lead:
        piece[lead_capacity - lead: lead_capacity] = buf[RANGE_LEAD - lead:].to(piece.device)
    return lead


def _peer(r: int, group) -> int:
    return dist.get_global_rank(group, r) if group is not None else r


def _send_int(v: int, dst: int, device, group):
    dist.send(torch.tensor([v], dtype=torch.int64, device=device), _peer(dst, group), group=group)


def _recv_int(src: int, device, group) -> int:
    t = torch.zeros(1, dtype=torch.int64, device=device)
    dist.recv(t, _peer(src, group), group=group)
    return int(t.item())


def splice_at_common_cut(fix_ends, fix_hashes, spec_ends, spec_hashes):
    """First cut of the re-synchronised chain that is also a cut of the speculative chain.  The chunk
    that ENDS at that cut still starts at a different place in the two chains, so it is taken from
    the re-synchronised one; everything after it is identical and comes from the speculative list.
    Returns (ends, hashes) or None if the chains share no cut."""
    common, fi, si = np.intersect1d(fix_ends, spec_ends, assume_unique=True, return_indices=True)
    if common.size == 0:
        return None
    i, j = int(fi[0]), int(si[0])
    return (np.concatenate([fix_ends[:i + 1], spec_ends[j + 1:]]),
            np.concatenate([fix_hashes[:i + 1], spec_hashes[j + 1:]]))


def hash_file_by_ranges(engine, piece, lead: int, file_len: int, begin: int, end: int, group=None,
                        comm_device: Optional[torch.device] = None, phases: Optional[dict] = None,
                        spec=None) -> RangeOutcome:
    """`piece` holds bytes [begin - lead, end) of the file (tensor / array / buffer the engine accepts,
    sliceable).  Collective over `group`: ranks must call it with consecutive ranges in rank order.
    `spec` = (chunk_ends, chunk_hashes, exit_cut) of the speculative pass when the caller has already run it
    (as one item of a batched call, `hash_concatenated`)."""
    import time as _time
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = comm_device or torch.device("cpu")
    rehashed = 0
    _t = [_time.perf_counter()]

    def _mark(name):
        if phases is not None:
            now = _time.perf_counter()
            phases[name] = phases.get(name, 0.0) + (now - _t[0]) * 1e3
            _t[0] = now

    # 2. speculative pass over the whole range (rank 0's entry is the true one: the file start)
    assumed = 0 if rank == 0 else begin
    if spec is not None:
        ends, hashes, exit_cut = spec
    elif end > begin or rank == 0:
        spec = engine.hash_range(piece, file_len, begin, end, assumed)
        ends, hashes, exit_cut = spec.chunk_ends, spec.chunk_hashes, spec.exit_cut
    else:                                          # empty range: owns no chunk, only passes the carry on
        ends, hashes, exit_cut = np.zeros(0, dtype=np.uint64), np.zeros((0, 32), dtype=np.uint8), assumed
    _mark("speculative_pass_ms")
    # 3. carry chain
    if rank > 0:
        true_entry = _recv_int(rank - 1, dev, group)
        if true_entry != assumed and end > begin:
            window = FIX_WINDOW
            while True:
                w_end = min(end, begin + window)
                fix = engine.hash_range(piece[: lead + (w_end - begin)], file_len, begin, w_end, true_entry)
                rehashed += lead + (w_end - begin)
                merged = splice_at_common_cut(fix.chunk_ends, fix.chunk_hashes, ends, hashes)
                if merged is not None:
                    ends, hashes = merged          # exit cut unchanged: the tail is the speculative chain
                    break
                if w_end == end:                   # chains never met inside this range: the re-hash is the answer
                    ends, hashes, exit_cut = fix.chunk_ends, fix.chunk_hashes, fix.exit_cut
                    break
                window *= 4
        elif end == begin:
            exit_cut = true_entry                  # empty range: pass the carry through
    if rank + 1 < world:
        _send_int(int(exit_cut), rank + 1, dev, group)
    out = RangeOutcome(ends, hashes, int(exit_cut), rehashed)
    _mark("carry_chain_ms")
    # 4. chunk lists to rank 0: one packed record array per rank (8-byte end + 32-byte hash), kept on the
    #    communication device so that rank 0 can hand the concatenation straight to the merkle combine
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    counts[rank] = ends.size
    dist.all_reduce(counts, group=group)
    counts = counts.cpu().numpy()

    def pack(e, h):
        rec = np.empty((e.size, 40), dtype=np.uint8)
        rec[:, :8] = np.ascontiguousarray(e, dtype="<u8").view(np.uint8).reshape(-1, 8)
        rec[:, 8:] = h
        return torch.from_numpy(rec).to(dev)

    if rank == 0:
        recs = [pack(ends, hashes)]
        for r in range(1, world):
            t = torch.empty(int(counts[r]), 40, dtype=torch.uint8, device=dev)
            if counts[r]:
                dist.recv(t, _peer(r, group), group=group)
            recs.append(t)
        rec = torch.cat(recs)
        e = rec[:, :8].contiguous().view(torch.int64).reshape(-1)
        h = rec[:, 8:].contiguous()
        sizes = torch.diff(e, prepend=torch.zeros(1, dtype=torch.int64, device=e.device))
        _mark("gather_ms")
        out.merkle_root, out.file_hash = engine.merkle_from_chunks(h, sizes)
        out.n_chunks_total = int(e.numel())
        _mark("merkle_ms")
    elif ends.size:
        dist.send(pack(ends, hashes), _peer(0, group), group=group)
        _mark("gather_ms")
    return out


# ------------------------------------------------------------- concatenated work (BASELINE config 4)
PIECE_HEADROOM = 1 << 18   # bytes kept free in front of a partial piece: >= RANGE_LEAD, keeps the piece 256-byte aligned


@dataclass
class ConcatLayout:
    """Where one rank keeps its pieces of a `plan_concatenated` plan inside one flat byte buffer."""
    pieces: List[Piece]
    offsets: List[int]        # buffer offset of the headroom (partial pieces) or of the first byte (whole files)
    headroom: List[int]       # PIECE_HEADROOM for a piece that does not start at byte 0 of its file, else 0
    nbytes: int               # buffer bytes needed

    def data_offset(self, k: int) -> int:
        return self.offsets[k] + self.headroom[k]


def layout_pieces(pieces: Sequence[Piece]) -> ConcatLayout:
    offs, heads, off = [], [], 0
    for p in pieces:
        head = PIECE_HEADROOM if p.begin > 0 else 0
        offs.append(off)
        heads.append(head)
        off = (off + head + (p.end - p.begin) + 255) & ~255
    return ConcatLayout(list(pieces), offs, heads, off + 256)


def shared_files(plan: Sequence[Sequence[Piece]]):
    """(owners: file -> ranks holding a piece of it, shared: files held by more than one rank, ascending)."""
    owners = {}
    for r, pieces in enumerate(plan):
        for p in pieces:
            owners.setdefault(p.file, []).append(r)
    return owners, sorted(f for f, rs in owners.items() if len(rs) > 1)


def hash_concatenated(engine, buf: torch.Tensor, lay: ConcatLayout, plan: Sequence[Sequence[Piece]], n_files: int,
                      groups: dict, comm_device: Optional[torch.device] = None, phases: Optional[dict] = None,
                      memory: int = 0):
    """One pass over a `plan_concatenated` job.  Collective: every rank calls it with its own buffer/layout.

    Everything a rank holds goes through ONE batched engine call: the files that lie inside its range as whole
    files, and its pieces of the files that straddle a range boundary as byte-range items hashed speculatively
    (halo exchanged first).  The shared files then finish the byte-range protocol (`hash_file_by_ranges`: carried
    cut, fix-up window, record gather, merkle combine) inside their own process groups `groups[file]` —
    even-numbered shared files first, then odd-numbered ones, so that no rank waits on a chain of neighbours.  Returns (file_hashes uint8 [n_files, 32] on every rank,
    max bytes any fix-up window re-hashed on this rank)."""
    import time as _time
    rank = dist.get_rank()
    dev = comm_device or torch.device("cpu")
    owners, shared = shared_files(plan)
    out = torch.zeros(n_files, 32, dtype=torch.int32)
    t0 = _time.perf_counter()
    # 1. halo exchange for every shared file this rank holds a piece of (even-numbered shared files first, then
    #    odd-numbered ones: no rank waits on a chain of neighbours)
    my_shared = {}   # file -> (piece index, lead)
    for parity in (0, 1):
        for j, f in enumerate(shared):
            if j % 2 != parity or rank not in owners[f]:
                continue
            k = next(i for i, p in enumerate(lay.pieces) if p.file == f)
            p, o, head = lay.pieces[k], lay.offsets[k], lay.headroom[k]
            ranges = [(q.begin, q.end) for r in owners[f] for q in plan[r] if q.file == f]
            t1 = _time.perf_counter()
            lead = exchange_lead(buf[o: o + head + (p.end - p.begin)], head, ranges, group=groups[f])
            if phases is not None:
                phases["halo_exchange_ms"] = phases.get("halo_exchange_ms", 0.0) + (_time.perf_counter() - t1) * 1e3
            my_shared[f] = (k, lead)
    # 2. ONE batched call for everything this rank holds: its whole files and, as byte-range items hashed from the
    #    assumption "a cut at my range begin", its pieces of the shared files.  The alignment padding / unused
    #    headroom between items is described to the engine as spacer files (their results are dropped).
    t0 = _time.perf_counter()
    for p in lay.pieces:
        if p.file_len == 0:   # an empty file has no bytes to lay out: ask the engine for its (all-zero) hash directly
            fr = engine.hash_buffers(buf[:0], np.zeros(2, dtype=np.uint64), memory=memory).files[0]
            out[p.file] = torch.frombuffer(bytearray(fr.file_hash), dtype=torch.uint8).to(torch.int32)
    items = []   # (start, end, spec, piece index or -1)
    for k, p in enumerate(lay.pieces):
        if p.file_len == 0 or p.end == p.begin:
            continue
        if p.whole:
            items.append((lay.offsets[k], lay.offsets[k] + p.file_len, (p.file_len, 0, p.file_len, 0), k))
        else:
            lead = my_shared[p.file][1]
            d0 = lay.offsets[k] + lay.headroom[k] - lead
            entry = 0 if p.begin == 0 else p.begin
            items.append((d0, d0 + lead + (p.end - p.begin), (p.file_len, p.begin, p.end, entry), k))
    spec_of = {}
    if items:
        items.sort()
        marks, specs, owner = [], [], []
        pos = items[0][0]
        for st, en, sp, k in items:
            if st > pos:   # spacer
                marks.append(pos); specs.append((st - pos, 0, st - pos, 0)); owner.append(-1)
            marks.append(st); specs.append(sp); owner.append(k)
            pos = en
        marks.append(pos)
        if all(k < 0 or lay.pieces[k].whole for k in owner):   # no byte-range item: the plain batched call
            cl = engine.hash_buffers(buf, np.asarray(marks, dtype=np.uint64), memory=memory)
            exits = None
        else:
            cl, exits = engine.hash_buffers_ranges(buf, np.asarray(marks, dtype=np.uint64), specs, memory=memory)
        for i, k in enumerate(owner):
            if k < 0:
                continue
            p = lay.pieces[k]
            if p.whole:
                out[p.file] = torch.frombuffer(bytearray(cl.files[i].file_hash), dtype=torch.uint8).to(torch.int32)
            else:
                e, h = cl.chunks_of(i)
                spec_of[p.file] = (e, h, exits[i])
    if phases is not None:
        phases["batched_pass_ms"] = phases.get("batched_pass_ms", 0.0) + (_time.perf_counter() - t0) * 1e3
    # 3. carry chain, fix-up window, record gather and merkle combine per shared file
    rehashed = 0
    for parity in (0, 1):
        for j, f in enumerate(shared):
            if j % 2 != parity or rank not in owners[f]:
                continue
            k, lead = my_shared[f]
            p, o, head = lay.pieces[k], lay.offsets[k], lay.headroom[k]
            view = buf[o: o + head + (p.end - p.begin)]
            res = hash_file_by_ranges(engine, view[head - lead:], lead, p.file_len, p.begin, p.end, group=groups[f],
                                      comm_device=dev, phases=phases, spec=spec_of.get(f))
            rehashed = max(rehashed, res.rehashed_bytes)
            if res.file_hash is not None:
                out[f] = torch.frombuffer(bytearray(res.file_hash), dtype=torch.uint8).to(torch.int32)
    # 4. every rank ends with every file hash (rows are non-zero on exactly one rank: a sum is a gather)
    t2 = _time.perf_counter()
    if dist.get_world_size() > 1:
        t = out.to(dev)
        dist.all_reduce(t)
        out = t.cpu()
    if phases is not None:
        phases["result_allreduce_ms"] = phases.get("result_allreduce_ms", 0.0) + (_time.perf_counter() - t2) * 1e3
    return out.to(torch.uint8).numpy(), rehashed
