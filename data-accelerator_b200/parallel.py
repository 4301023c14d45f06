"""Multi-GPU plumbing for one 8xB200 box: one process per GPU (torchrun), torch.distributed for the
rendezvous and the all-to-all; the partition/pack/unpack work is done by libdxgpu's K5 kernels.

Stands in for Spark's hash-partitioned shuffle between partial and final aggregation (the Exchange
behind every GROUP BY executed at CommonProcessorFactory.scala:271): rows are sent to the rank that
owns hash64(group key) % world; that rank's flow holds the key's window cache and state tables.

`exchange_counts` / `exchange_records` are backend-agnostic (nccl on GPUs; gloo in the CPU tests,
where the buffers are numpy/torch CPU tensors produced by the oracle's partitioner).
"""
import ctypes as C

import numpy as np

from . import _native as N
from .device import Table

vp, i64, i32, cp = C.c_void_p, C.c_int64, C.c_int, C.c_char_p
P = C.POINTER
_bound = False


def _lib():
    global _bound
    L = N.lib()
    if not _bound:
        L.dx_flow_project.restype, L.dx_flow_project.argtypes = i32, [vp, vp, i64, vp, i64, i32, i64, P(vp)]
        L.dx_flow_process_table.restype, L.dx_flow_process_table.argtypes = i32, [vp, cp, vp, i64, i64, i64, P(vp)]
        L.dx_table_shuffle_pack.restype = i32
        L.dx_table_shuffle_pack.argtypes = [vp, P(cp), i32, P(cp), i32, i32, vp, P(i64), P(i64)]
        L.dx_table_from_packed.restype = i32
        L.dx_table_from_packed.argtypes = [vp, P(cp), P(i32), i32, vp, i64, P(vp)]
        _bound = True
    return L


def project_device(proc, ptr, n_bytes, n_events, now_ms):
    """Decode + projection of a device-resident batch -> Table (no windows / SQL)."""
    L = _lib()
    t = C.c_void_p()
    proc.ctx.check(L.dx_flow_project(proc.h, C.c_void_p(ptr), n_bytes, None, n_events, 1, now_ms, C.byref(t)))
    return Table(proc.ctx, t)


def process_table(proc, table, batch_ms, interval_ms, now_ms=None, namespace=""):
    """route() on a projected table (e.g. rows received from the shuffle)."""
    from .processor import BatchResult
    L = _lib()
    r = C.c_void_p()
    proc.ctx.check(L.dx_flow_process_table(proc.h, namespace.encode(), table.h, batch_ms, interval_ms,
                                           batch_ms if now_ms is None else now_ms, C.byref(r)))
    return BatchResult(proc.ctx, r)


def pack_by_key(table, cols, key_cols, world, out_ptr):
    """K5: fixed records grouped by owner rank into out_ptr (device).  -> (counts[world], rec_bytes)."""
    L = _lib()
    c_cols = (cp * len(cols))(*[c.encode() for c in cols])
    c_keys = (cp * len(key_cols))(*[c.encode() for c in key_cols])
    counts = (i64 * world)()
    rec = i64()
    table.ctx.check(L.dx_table_shuffle_pack(table.h, c_cols, len(cols), c_keys, len(key_cols), world, C.c_void_p(out_ptr), counts, C.byref(rec)))
    return np.array(list(counts), dtype=np.int64), rec.value


def table_from_packed(ctx, names, types, ptr, n_rows):
    L = _lib()
    c_names = (cp * len(names))(*[n.encode() for n in names])
    c_types = (i32 * len(types))(*types)
    t = C.c_void_p()
    ctx.check(L.dx_table_from_packed(ctx.h, c_names, c_types, len(names), C.c_void_p(ptr), n_rows, C.byref(t)))
    return Table(ctx, t)


# ---- backend-agnostic exchange (torch.distributed) ---------------------------------------------------
def exchange_counts(send_counts, group=None):
    """send_counts[r] = records this rank sends to rank r  ->  recv_counts[r] = records it receives from r."""
    import torch
    import torch.distributed as dist
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    s = torch.as_tensor(np.asarray(send_counts, dtype=np.int64), device=dev)
    r = torch.empty_like(s)
    dist.all_to_all_single(r, s, group=group)
    return r.cpu().numpy()


def exchange_records(send_buf, send_counts, recv_counts, rec_words, group=None):
    """All-to-all-v of fixed-size records.  send_buf: 1-D int64 tensor laid out by destination rank
    (send_counts[r] records of rec_words words each); returns the received 1-D int64 tensor."""
    import torch
    import torch.distributed as dist
    recv = torch.empty(int(np.sum(recv_counts)) * rec_words, dtype=torch.int64, device=send_buf.device)
    in_split = [int(c) * rec_words for c in send_counts]
    out_split = [int(c) * rec_words for c in recv_counts]
    dist.all_to_all_single(recv, send_buf, output_split_sizes=out_split, input_split_sizes=in_split, group=group)
    return recv


def shuffle_table(table, cols, types, key_cols, group=None):
    """Full K5 -> all-to-all -> K5 round trip for one table on the current rank (GPU, nccl)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = table.num_rows
    rec_words = len(cols) + 1
    send = torch.empty(max(1, n) * rec_words, dtype=torch.int64, device="cuda")
    counts, rec_bytes = pack_by_key(table, cols, key_cols, world, send.data_ptr())
    assert rec_bytes == rec_words * 8
    torch.cuda.synchronize()
    recv_counts = exchange_counts(counts, group)
    recv = exchange_records(send[: n * rec_words], counts, recv_counts, rec_words, group)
    torch.cuda.synchronize()
    n_recv = int(recv_counts.sum())
    out = table_from_packed(table.ctx, cols, types, recv.data_ptr() if n_recv else 0, n_recv)
    return out, {"sent_records": int(counts.sum()), "received_records": n_recv, "rec_bytes": rec_bytes,
                 "wire_bytes_out": int((counts.sum() - counts[dist.get_rank(group)]) * rec_bytes)}


# ---- pack fused with the exchange: records written straight into the owners' buffers over NVLink ----
def count_by_key(table, cols, key_cols, world):
    """K5 count only: rows of `table` per owner rank."""
    L = N.lib()
    L.dx_table_shuffle_count.restype, L.dx_table_shuffle_count.argtypes = i32, [vp, P(cp), i32, P(cp), i32, i32, P(i64)]
    c_cols = (cp * len(cols))(*[c.encode() for c in cols])
    c_keys = (cp * len(key_cols))(*[c.encode() for c in key_cols])
    counts = (i64 * world)()
    table.ctx.check(L.dx_table_shuffle_count(table.h, c_cols, len(cols), c_keys, len(key_cols), world, counts))
    return np.array(list(counts), dtype=np.int64)


def put_by_key(table, cols, key_cols, world, peer_ptrs, first_record):
    """K5 fused put: every record goes to peer_ptrs[owner] at record index first_record[owner] + (position among this
    rank's records for that owner)."""
    L = N.lib()
    L.dx_table_shuffle_put.restype, L.dx_table_shuffle_put.argtypes = i32, [vp, P(cp), i32, P(cp), i32, i32, P(vp), P(i64)]
    c_cols = (cp * len(cols))(*[c.encode() for c in cols])
    c_keys = (cp * len(key_cols))(*[c.encode() for c in key_cols])
    ptrs = (vp * world)(*[C.c_void_p(int(p)) for p in peer_ptrs])
    first = (i64 * world)(*[int(x) for x in first_record])
    table.ctx.check(L.dx_table_shuffle_put(table.h, c_cols, len(cols), c_keys, len(key_cols), world, ptrs, first))


def put_blocks_by_key(table, cols, key_cols, world, peer_ptrs, first_record, block_records):
    """K5 fused put into fixed per-sender blocks (no count exchange before it) -> records written per owner."""
    L = N.lib()
    L.dx_table_shuffle_put_blocks.restype = i32
    L.dx_table_shuffle_put_blocks.argtypes = [vp, P(cp), i32, P(cp), i32, i32, P(vp), P(i64), P(i64), P(i64)]
    c_cols = (cp * len(cols))(*[c.encode() for c in cols])
    c_keys = (cp * len(key_cols))(*[c.encode() for c in key_cols])
    ptrs = (vp * world)(*[C.c_void_p(int(p)) for p in peer_ptrs])
    first = (i64 * world)(*[int(x) for x in first_record])
    blocks = (i64 * world)(*[int(x) for x in block_records])
    sent = (i64 * world)()
    table.ctx.check(L.dx_table_shuffle_put_blocks(table.h, c_cols, len(cols), c_keys, len(key_cols), world, ptrs, first, blocks, sent))
    return np.array(list(sent), dtype=np.int64)


def table_from_packed_blocks(ctx, names, types, ptr, block_first, block_count):
    """K5 unpack of records that lie in one run per sender (block_first / block_count in records)."""
    L = _lib()
    L.dx_table_from_packed_blocks.restype = i32
    L.dx_table_from_packed_blocks.argtypes = [vp, P(cp), P(i32), i32, vp, i32, P(i64), P(i64), P(vp)]
    c_names = (cp * len(names))(*[n.encode() for n in names])
    c_types = (i32 * len(types))(*types)
    nb = len(block_first)
    bf = (i64 * nb)(*[int(x) for x in block_first])
    bc = (i64 * nb)(*[int(x) for x in block_count])
    t = C.c_void_p()
    ctx.check(L.dx_table_from_packed_blocks(ctx.h, c_names, c_types, len(names), C.c_void_p(ptr), nb, bf, bc, C.byref(t)))
    return Table(ctx, t)


def block_offsets(M, rank):
    """M[s][d] = records rank s sends to rank d  ->  (first[d]: where rank `rank`'s block starts in rank d's receive buffer —
    blocks are laid out in sender order — and recv_total[d]: records rank d receives in all)."""
    M = np.asarray(M, dtype=np.int64)
    first = M[:rank].sum(axis=0) if rank else np.zeros(M.shape[1], dtype=np.int64)
    return first, M.sum(axis=0)


class PeerExchange:
    """Receive buffers of all ranks of one box, mapped into every process (CUDA IPC over NVLink), and the bookkeeping of
    one fused exchange per micro-batch: put into per-sender blocks -> all_gather of the counts -> local unpack (exchange_blocks),
    or counts -> all_gather -> block offsets -> put -> barrier -> local unpack (exchange: blocks laid out back to back).
    torch.distributed carries only the small integers and the barrier; the records never pass through a collective."""

    def __init__(self, ctx, capacity_records, rec_words, group=None):
        import torch
        import torch.distributed as dist
        self.ctx, self.group, self.rec_words, self.capacity = ctx, group, rec_words, int(capacity_records)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        L = N.lib()
        L.dx_ipc_export.restype, L.dx_ipc_export.argtypes = i32, [vp, vp, C.c_char_p]
        L.dx_ipc_open.restype, L.dx_ipc_open.argtypes = i32, [vp, C.c_char_p, P(vp)]
        L.dx_ipc_close.restype, L.dx_ipc_close.argtypes = i32, [vp, vp]
        self.L = L
        # two receive buffers, used by alternate batches (exchange_blocks: a fast rank's next put must not land in the buffer a
        # slow rank is still unpacking)
        self.locals = [ctx.alloc(self.capacity * rec_words * 8) for _ in range(2)]
        self.local = self.locals[0]
        self.batch_no = 0
        raw = bytearray()
        for buf in self.locals:
            h = C.create_string_buffer(64)
            ctx.check(L.dx_ipc_export(ctx.h, C.c_void_p(buf), h))
            raw += h.raw
        mine = torch.frombuffer(raw, dtype=torch.uint8).cuda()
        allh = [torch.empty(128, dtype=torch.uint8, device="cuda") for _ in range(self.world)]
        dist.all_gather(allh, mine, group=group)
        self.ptrs2 = [[], []]
        for r in range(self.world):
            hb = allh[r].cpu().numpy().tobytes()
            for k in range(2):
                if r == self.rank:
                    self.ptrs2[k].append(self.locals[k])
                    continue
                p = C.c_void_p()
                ctx.check(L.dx_ipc_open(ctx.h, hb[64 * k: 64 * k + 64], C.byref(p)))
                self.ptrs2[k].append(p.value)
        self.ptrs = self.ptrs2[0]

    def exchange(self, table, cols, key_cols):
        """-> (device pointer of this rank's received records, number of records, counts this rank sent)."""
        import torch
        import torch.distributed as dist
        counts = count_by_key(table, cols, key_cols, self.world)
        mine = torch.as_tensor(counts, device="cuda")
        allc = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(allc, mine, group=self.group)          # also orders this batch after every rank's previous unpack
        M = torch.stack(allc).cpu().numpy()                    # M[s][d] = records rank s sends to rank d
        first, recv_total = block_offsets(M, self.rank)
        if int(recv_total.max()) > self.capacity:
            raise RuntimeError("PeerExchange: %d records for one rank exceed the buffer capacity %d" % (int(recv_total.max()), self.capacity))
        put_by_key(table, cols, key_cols, self.world, self.ptrs, first)
        dist.barrier(group=self.group)                         # every rank's puts have landed
        torch.cuda.synchronize()                               # (the barrier runs on torch's stream, the unpack on the library's)
        return self.local, int(recv_total[self.rank]), counts

    def exchange_blocks(self, table, cols, key_cols):
        """The exchange without a count round before it: every sender owns a fixed block (capacity / world records) in every
        owner's buffer, puts straight away, and ONE all-gather of the per-owner counts afterwards both tells the owners what
        arrived and that every put has landed.  -> (receive buffer, first record of each sender's block, records in each
        block, counts this rank sent); unpack with table_from_packed_blocks."""
        import torch
        import torch.distributed as dist
        par = self.batch_no & 1
        self.batch_no += 1
        block = self.capacity // self.world
        sent = put_blocks_by_key(table, cols, key_cols, self.world, self.ptrs2[par], [self.rank * block] * self.world, [block] * self.world)
        mine = torch.as_tensor(sent, device="cuda")
        allc = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(allc, mine, group=self.group)          # a rank enters only after its put kernel has completed
        M = torch.stack(allc).cpu().numpy()                    # M[s][d] = records rank s sent to rank d
        return self.locals[par], [s_ * block for s_ in range(self.world)], M[:, self.rank].copy(), sent

    def close(self):
        for k in range(2):
            for r, p in enumerate(self.ptrs2[k]):
                if r != self.rank and p:
                    self.L.dx_ipc_close(self.ctx.h, C.c_void_p(p))
        for buf in self.locals:
            self.ctx.free(buf)
        self.ptrs = []
        self.ptrs2 = [[], []]
