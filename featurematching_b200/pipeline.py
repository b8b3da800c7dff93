"""Sync-free matching head for a batch of frame pairs, and the cross-GPU match-list gather.

`MatchHead` chains the three drop-in modules of modules.py the way network/net.py:75-83 does
(coarse match -> window gather -> [fine LocalFeatureTransformer, optional] -> fine match), but keeps the match count on the device: every buffer is sized for `capacity` matches
and the kernels read the count from the coarse status word, so a batch costs one host
synchronisation (at the caller's choosing) instead of the reference's one per stage
(torch.where at coarse_matching_new.py:109, the shape test at fine_preprocess.py:37).

`gather_matches` is the one cross-GPU step: ranks own disjoint frame pairs; their packed match
records go to the consumer rank with one grouped send / receive (or to every rank with one padded
all_gather) -- NCCL over NVLink on GPUs, gloo in the CPU tests; the analogue of the reference's pickled
gather / all_gather (utils/comm.py:141-219).
"""
import numpy as np
import torch

from . import _lib, modules

RECORD_DTYPE = np.dtype([("pair", "<i4"), ("i", "<i4"), ("j", "<i4"), ("conf", "<f4"),
                         ("x0", "<f4"), ("y0", "<f4"), ("std0", "<f4"),
                         ("x1", "<f4"), ("y1", "<f4"), ("std1", "<f4")])
assert RECORD_DTYPE.itemsize == _lib.RECORD_BYTES


class HeadStatusError(_lib.FmError):
    """A device-reported condition of the coarse stage (status word of fm_coarse_match)."""

    def __init__(self, flags, status):
        super().__init__(f"coarse status flags {flags:#x} (cand overflow=1, output overflow=2, "
                         f"tie overflow=4, range=8); status={status}")
        self.flags, self.status = flags, status


class HeadResult:
    """Device-side outputs of one MatchHead.run call (padded to `capacity`)."""
    __slots__ = ("coarse", "win0", "win1", "mkpts0_f", "mkpts1_f", "records", "capacity")

    @property
    def count(self):
        return self.coarse.status[_lib.ST_COUNT:_lib.ST_COUNT + 1]


class MatchHead:
    def __init__(self, config, device, window_dtype=torch.float32, fine_transformer=None):
        self.device = torch.device(device)
        self.coarse_matching = modules.CoarseMatching(config['match_coarse']).eval()
        self.fine_preprocess = modules.FinePreprocess(config).eval().to(self.device)
        self.fine_preprocess.out_dtype = window_dtype
        self.fine_matching = modules.FineMatching(config['fine']).eval().to(self.device)
        # fine LocalFeatureTransformer (network/net.py:79-80): True builds the B200 module from
        # config['fine'] (random init, load a state dict into head.fine_transformer), a
        # modules.LocalFeatureTransformer is used as given (both stay sync-free), any other
        # callable is treated as a foreign PyTorch module (needs exact-length inputs: one sync)
        if fine_transformer is True:
            fine_transformer = modules.LocalFeatureTransformer(config['fine'])
        if isinstance(fine_transformer, modules.LocalFeatureTransformer):
            fine_transformer = fine_transformer.eval().to(self.device)
        self.fine_transformer = fine_transformer
        self._res = None
        self._host = {}

    def load_params(self, p):
        """p: dict with down_w/down_b/merge_w/merge_b/mix_w0/mix_b0/mix_w1/mix_b1 (synth.hot_path_params)."""
        with torch.no_grad():
            pre, fin = self.fine_preprocess, self.fine_matching
            if pre.cat_c_feat:
                pre.down_proj.weight.copy_(p["down_w"]); pre.down_proj.bias.copy_(p["down_b"])
                pre.merge_feat.weight.copy_(p["merge_w"]); pre.merge_feat.bias.copy_(p["merge_b"])
            fin.mix_feat_0.weight.copy_(p["mix_w0"].reshape(1, -1)); fin.mix_feat_0.bias.copy_(p["mix_b0"].reshape(1))
            fin.mix_feat_1.weight.copy_(p["mix_w1"].reshape(1, -1)); fin.mix_feat_1.bias.copy_(p["mix_b1"].reshape(1))
        self._params = None

    @staticmethod
    def default_capacity(B, hw0_c, hw1_c, border_rm):
        """Mutual nearest neighbours of interior cells: at most one per interior cell of the
        smaller grid (exact ties aside, which the status flags report)."""
        n0 = max(hw0_c[0] - 2 * border_rm, 0) * max(hw0_c[1] - 2 * border_rm, 0)
        n1 = max(hw1_c[0] - 2 * border_rm, 0) * max(hw1_c[1] - 2 * border_rm, 0)
        return max(B * min(n0, n1), 1)

    def _buffers(self, capacity, WW, Cf):
        r = self._res
        if r is not None and r.capacity >= capacity:
            return r
        dev = self.device
        r = HeadResult()
        r.capacity = capacity
        r.coarse = None
        wd = self.fine_preprocess.out_dtype
        r.win0 = torch.empty(capacity, WW, Cf, dtype=wd, device=dev)
        r.win1 = torch.empty(capacity, WW, Cf, dtype=wd, device=dev)
        r.mkpts0_f = torch.empty(capacity, 3, dtype=torch.float32, device=dev)
        r.mkpts1_f = torch.empty(capacity, 3, dtype=torch.float32, device=dev)
        r.records = torch.empty((capacity + 1) * _lib.RECORD_BYTES, dtype=torch.uint8, device=dev)
        self._res = r
        return r

    def run(self, feat_c0, feat_c1, feat_f0, feat_f1, hw0_i, hw0_c, hw1_c, hw0_f, capacity=None,
            pair_offset=0, pack=True, append=None):
        """Enqueues the whole head on the current stream; no host synchronisation.
        feat_c*: [B, N, C] bf16/fp32; feat_f*: [B, Cf, Hf, Wf] (channels-last bf16 is read in place)."""
        B = feat_c0.shape[0]
        cm, pre, fin = self.coarse_matching, self.fine_preprocess, self.fine_matching
        if capacity is None:
            capacity = self.default_capacity(B, hw0_c, hw1_c, cm.border_rm)
        WW, Cf = pre.W * pre.W, feat_f0.shape[1]
        r = self._buffers(capacity, WW, Cf)
        with torch.no_grad():
            r.coarse = cm.match_raw(feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, capacity=r.capacity, out=r.coarse)
            c = r.coarse
            count = c.status   # status[0] = match count
            pre.gather(feat_f0, feat_f1, feat_c0, feat_c1, c.b_ids, c.i_ids, c.j_ids, hw0_c, hw1_c,
                       hw0_f, count=count, capacity=r.capacity, out=(r.win0, r.win1))
            w0, w1 = r.win0, r.win1
            if isinstance(self.fine_transformer, modules.LocalFeatureTransformer):
                if r.win0.dtype != torch.float32:
                    raise _lib.FmError("the fine transformer kernel takes fp32 windows")
                self.fine_transformer.transform(r.win0, r.win1, count=count, out=(r.win0, r.win1))
            elif self.fine_transformer is not None:
                # foreign PyTorch module; needs exact-length inputs
                M = int(count[0].item())
                w0, w1 = self.fine_transformer(r.win0[:M], r.win1[:M])
            scale = float(hw0_i[0]) / float(hw0_f[0])
            fin.refine(w0, w1, c.mkpts0_c, c.mkpts1_c, scale, count=count,
                       out=(r.mkpts0_f, r.mkpts1_f))
            if append is not None:
                # chunked batch: records go behind those of the earlier chunks (fm_pack_matches_append)
                counts, index, total_capacity, records = append
                lib = _lib.load()
                with torch.cuda.device(self.device):
                    rc = lib.fm_pack_matches_append(_lib.ptr(c.b_ids), _lib.ptr(c.i_ids), _lib.ptr(c.j_ids),
                                                    _lib.ptr(c.mconf), _lib.ptr(r.mkpts0_f), _lib.ptr(r.mkpts1_f),
                                                    _lib.ptr(count), r.capacity, int(pair_offset), _lib.ptr(counts),
                                                    int(index), int(total_capacity), _lib.ptr(records),
                                                    _lib.stream_ptr(self.device))
                _lib.check(rc, "fm_pack_matches_append")
            elif pack:
                lib = _lib.load()
                with torch.cuda.device(self.device):
                    rc = lib.fm_pack_matches(_lib.ptr(c.b_ids), _lib.ptr(c.i_ids), _lib.ptr(c.j_ids),
                                             _lib.ptr(c.mconf), _lib.ptr(r.mkpts0_f), _lib.ptr(r.mkpts1_f),
                                             _lib.ptr(count), r.capacity, int(pair_offset),
                                             _lib.ptr(r.records), _lib.stream_ptr(self.device))
                _lib.check(rc, "fm_pack_matches")
        return r

    def check_status(self, r):
        """One small synchronising copy; raises on device-reported overflow / range loss."""
        st = self.coarse_matching.read_status(r.coarse)
        flags = st[_lib.ST_FLAGS]
        if flags:
            raise HeadStatusError(flags, st)
        return st[_lib.ST_COUNT]

    def run_checked(self, *args, **kwargs):
        """run() + check_status() that repairs what can be repaired instead of raising: candidate-buffer
        overflow (even after the on-device exact extraction pass) grows cand_per_row and runs again, output
        overflow grows the capacity.  Returns (result, match count); one host synchronisation per attempt."""
        while True:
            r = self.run(*args, **kwargs)
            try:
                return r, self.check_status(r)
            except HeadStatusError as e:
                cm = self.coarse_matching
                if e.flags & _lib.FLAG_CAND_OVERFLOW and cm.cand_per_row < 256:
                    cm.cand_per_row = min(256, cm.cand_per_row * 4)
                elif e.flags & _lib.FLAG_OUT_OVERFLOW and not (e.flags & ~_lib.FLAG_OUT_OVERFLOW):
                    kwargs["capacity"] = r.capacity * 4
                else:
                    raise

    # ---------------------------------------------------------------- host-buffer entry point
    def _host_state(self, shapes, chunk, capacity):
        """Double-buffered device staging for `chunk` pairs, pinned landing buffers for the
        records / status words of every chunk, the copy stream and its events."""
        key = (tuple(shapes), chunk, capacity)
        st = self._host.get("state")
        if st is not None and st["key"] == key:
            return st
        dev = self.device
        st = {"key": key, "copy_stream": torch.cuda.Stream(device=dev), "stage": [], "free": [], "ready": []}
        for _ in range(2):
            bufs = []
            for shape, dtype, strides in shapes:
                t = torch.empty_strided((chunk,) + tuple(shape[1:]), strides, dtype=dtype, device=dev)
                bufs.append(t)
            st["stage"].append(bufs)
            st["free"].append(torch.cuda.Event())
            st["ready"].append(torch.cuda.Event())
        self._host["state"] = st
        return st

    def match_host_submit(self, feat_c0, feat_c1, feat_f0, feat_f1, hw0_i, hw0_c, hw1_c, hw0_f, pair_offset=0,
                          chunks=16):
        """First half of match_host: enqueues the H2D copies (copy stream), the head (current stream) and the D2H
        copies of the records / status words, and returns a ticket WITHOUT synchronising.  Up to two tickets may
        be outstanding (two pinned landing sets): submitting call k + 1 before collecting call k keeps the PCIe link
        busy across calls -- call k + 1's first groups cross while call k's last group still runs."""
        import time
        t_start = time.perf_counter()
        B = feat_c0.shape[0]
        nch = max(1, min(int(chunks), B, _lib.MAX_CHUNKS))
        chunk = (B + nch - 1) // nch
        nch = (B + chunk - 1) // chunk
        cm = self.coarse_matching
        capacity = self.default_capacity(chunk, hw0_c, hw1_c, cm.border_rm)
        ins = (feat_c0, feat_c1, feat_f0, feat_f1)
        st = self._host_state([(t.shape, t.dtype, t.stride()) for t in ins], chunk, capacity)
        total_capacity = nch * capacity
        rec_bytes = (total_capacity + 1) * _lib.RECORD_BYTES
        land = self._host.get("land")
        if land is None or land["rec_bytes"] != rec_bytes or land["nch"] != nch:
            land = {"rec_bytes": rec_bytes, "nch": nch, "turn": 0, "busy": [False, False],
                    "host": [(torch.empty(rec_bytes, dtype=torch.uint8).pin_memory(),
                              torch.empty(nch, _lib.STATUS_WORDS, dtype=torch.int32).pin_memory(),
                              torch.cuda.Event()) for _ in range(2)],
                    "rec_dev": torch.empty(rec_bytes, dtype=torch.uint8, device=self.device),
                    "counts": torch.zeros(_lib.MAX_CHUNKS, dtype=torch.int32, device=self.device)}
            self._host["land"] = land
        idx = land["turn"]
        if land["busy"][idx]:
            raise _lib.FmError("match_host_submit: two calls are already outstanding; collect one with match_host_result")
        land["turn"] ^= 1
        land["busy"][idx] = True
        rec_host, status_host, done = land["host"][idx]
        rec_dev, counts_dev = land["rec_dev"], land["counts"]
        cur = torch.cuda.current_stream(self.device)
        cs = st["copy_stream"]
        used = st.setdefault("used", [False, False])
        h2d = 0
        for k in range(nch):
            lo, hi = k * chunk, min((k + 1) * chunk, B)
            slot = k & 1
            with torch.cuda.stream(cs):
                if used[slot]:
                    cs.wait_event(st["free"][slot])   # the group (of this or of the previous call) that read this slot
                for d, t in zip(st["stage"][slot], ins):
                    d[:hi - lo].copy_(t[lo:hi], non_blocking=True)
                    h2d += (hi - lo) * t[0].numel() * t.element_size()
                st["ready"][slot].record(cs)
            cur.wait_event(st["ready"][slot])
            dev_in = [d[:hi - lo] for d in st["stage"][slot]]
            r = self.run(*dev_in, hw0_i, hw0_c, hw1_c, hw0_f, capacity=capacity, pair_offset=pair_offset + lo,
                         pack=False, append=(counts_dev, k, total_capacity, rec_dev))
            status_host[k].copy_(r.coarse.status, non_blocking=True)
            st["free"][slot].record(cur)
            used[slot] = True
        rec_host.copy_(rec_dev, non_blocking=True)
        done.record(cur)
        return {"idx": idx, "h2d": h2d, "d2h": rec_bytes + nch * _lib.STATUS_WORDS * 4, "t_start": t_start,
                "t_enq": time.perf_counter()}

    def match_host_result(self, ticket):
        """Second half: waits for the ticket's device work (one host synchronisation on its event) and returns
        (records, bytes host->device, bytes device->host) like match_host.  The records VIEW the ticket's pinned
        landing buffer, which the call after next reuses -- copy them to keep them."""
        import time
        land = self._host["land"]
        rec_host, status_host, done = land["host"][ticket["idx"]]
        done.synchronize()
        t_sync = time.perf_counter()
        land["busy"][ticket["idx"]] = False
        for k, stw in enumerate(status_host.tolist()):
            if stw[_lib.ST_FLAGS]:
                raise _lib.FmError(f"coarse status flags {stw[_lib.ST_FLAGS]:#x} in chunk {k} (cand overflow=1, "
                                   f"output overflow=2, tie overflow=4, range=8); status={stw}")
        a = rec_host.numpy().view(RECORD_DTYPE)
        rec = a[1:1 + int(a[0]["pair"])]
        # host-side split of the call (ms): enqueue, wait for the device, unpack
        self.last_timing = ((ticket["t_enq"] - ticket["t_start"]) * 1e3, (t_sync - ticket["t_enq"]) * 1e3,
                            (time.perf_counter() - t_sync) * 1e3)
        return rec, ticket["h2d"], ticket["d2h"]

    def match_host(self, feat_c0, feat_c1, feat_f0, feat_f1, hw0_i, hw0_c, hw1_c, hw0_f, pair_offset=0,
                   chunks=16):
        """End-to-end call on HOST tensors (pinned for full copy bandwidth).  The batch is cut into
        `chunks` groups of pairs; group k + 1 crosses PCIe on a copy stream while group k runs the
        head; the groups' packed records are appended on the device (fm_pack_matches_append) and
        come back, with the status words, in asynchronous copies.  One host synchronisation per
        call.  Returns (records as a numpy structured array (RECORD_DTYPE) that VIEWS a pinned
        landing buffer reused by later calls -- copy it to keep it --, bytes host->device, bytes
        device->host).  = match_host_result(match_host_submit(...)); a caller with a stream of batches
        submits batch k + 1 before collecting batch k."""
        return self.match_host_result(self.match_host_submit(feat_c0, feat_c1, feat_f0, feat_f1, hw0_i, hw0_c, hw1_c,
                                                              hw0_f, pair_offset=pair_offset, chunks=chunks))


def unpack_records(buf, capacity, world):
    """[world * (capacity + 1) * 40] uint8 (host) -> list of per-rank structured arrays."""
    a = buf.numpy().view(RECORD_DTYPE).reshape(world, capacity + 1)
    return [a[r, 1:1 + int(a[r, 0]["pair"])] for r in range(world)]


def gather_matches(records, capacity, group=None, dst=None):
    """The one cross-GPU step: ranks own disjoint frame pairs and exchange their packed match records.

    dst = None: every rank receives every rank's records (one padded all_gather).
    dst = r:    only the consumer rank r receives them (dist.gather = grouped ncclSend / ncclRecv: the seven
                other ranks of a box run a 5.8 MB send instead of an 8 x 5.8 MB all-gather kernel next to
                their persistent 148-CTA kernels) -- the role of the reference's `gather`, utils/comm.py:179-219;
                returns None on the other ranks.
    No size exchange and no host synchronisation either way (the count travels in each rank's header record).
    Every rank must call it, also with zero matches.  `records`: uint8 [(capacity + 1) * 40] on the rank's
    device (CPU tensors with the gloo backend).  Returns uint8 [world * (capacity + 1) * 40] or None."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = (capacity + 1) * _lib.RECORD_BYTES
    mine = records[:n].contiguous()
    if dst is None:
        out = torch.empty(world * n, dtype=torch.uint8, device=records.device)
        dist.all_gather_into_tensor(out, mine, group=group)
        return out
    if dist.get_rank(group) == dst:
        out = torch.empty(world * n, dtype=torch.uint8, device=records.device)
        dist.gather(mine, [out[r * n:(r + 1) * n] for r in range(world)], dst=dst, group=group)
        return out
    dist.gather(mine, None, dst=dst, group=group)
    return None


class OverlappedGather:
    """The match-list exchange on a side stream, placed under the NEXT step's HBM-bound fine kernels rather than
    under its tensor-core kernel ("gather once per batch, overlap with the next batch's compute"):

        g = OverlappedGather(device, dst=0)        # dst=None: all_gather to every rank
        for batch in batches:
            coarse stage of this batch             # the persistent 148-CTA kernel keeps every SM to itself
            g.launch()                             # exchange of the PREVIOUS batch starts here, behind it
            gather / fine transformer / fine match
            pack records into r.records
            g.stage(r.records, r.capacity)         # copies them into one of two staging slots; nothing is sent yet
        out = g.finish()                           # sends the last batch, joins the side stream

    No host synchronisation anywhere; `out` (on the consumer rank, None elsewhere when dst is set) belongs to the
    most recent exchange and is valid once the current stream has passed finish()."""

    def __init__(self, device, group=None, dst=None):
        self.device = torch.device(device)
        self.group, self.dst = group, dst
        self.side = torch.cuda.Stream(device=self.device)
        # two staging slots: stage() copies the caller's records (a few MB, device to device), so the caller's
        # buffer is free at once and an exchange has TWO steps to complete before its slot is needed again -- with
        # one buffer the pack of step k + 1 waited for exchange k, and a send cannot finish before the consumer
        # rank has posted the matching receive: the ranks ran in lockstep plus skew (r02: 8.8 vs 7.0 ms per step
        # on the sending rank of a 2-rank sequence job)
        self.slots = [None, None]
        self.packed = [torch.cuda.Event(), torch.cuda.Event()]
        self.done = [torch.cuda.Event(), torch.cuda.Event()]
        self.busy = [False, False]
        self.gate = torch.cuda.Event()
        self.turn = 0
        self.staged = None
        self.out = None

    def stage(self, records, capacity):
        n = (capacity + 1) * _lib.RECORD_BYTES
        s = self.turn
        self.turn ^= 1
        cur = torch.cuda.current_stream(self.device)
        if self.busy[s]:
            cur.wait_event(self.done[s])      # the exchange that used this slot two steps ago
            self.busy[s] = False
        if self.slots[s] is None or self.slots[s].numel() != n:
            self.slots[s] = torch.empty(n, dtype=torch.uint8, device=self.device)
        self.slots[s].copy_(records[:n], non_blocking=True)
        self.packed[s].record(cur)
        if self.staged is not None:           # stage() twice without launch(): send the older one now
            self.launch()
        self.staged = (s, capacity)

    def launch(self):
        if self.staged is None:
            return self.out
        s, capacity = self.staged
        self.staged = None
        self.gate.record(torch.cuda.current_stream(self.device))   # behind whatever the caller enqueued so far
        self.side.wait_event(self.packed[s])
        self.side.wait_event(self.gate)
        with torch.cuda.stream(self.side):
            self.out = gather_matches(self.slots[s], capacity, self.group, self.dst)
            self.done[s].record(self.side)
        self.busy[s] = True
        return self.out

    def finish(self):
        """Sends what is staged and joins the side stream into the current one (device-side)."""
        self.launch()
        cur = torch.cuda.current_stream(self.device)
        for s in range(2):
            if self.busy[s]:
                cur.wait_event(self.done[s])
                self.busy[s] = False
        return self.out


def shard_pairs(num_pairs, rank, world):
    """Contiguous block partition of frame pairs over ranks: [begin, end)."""
    base, rem = divmod(num_pairs, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_sequence(num_frames, rank, world):
    """Consecutive-frame matching of a sequence (pairs (k, k + 1), the pattern of
    datasets/scared_new2.py:76-83): rank owns pairs [begin, end) and therefore frames
    [begin, end] -- a one-frame halo, so no pair straddles two ranks."""
    begin, end = shard_pairs(max(num_frames - 1, 0), rank, world)
    return begin, end, (begin, end + 1 if end > begin else begin)
