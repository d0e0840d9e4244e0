"""Multi-GPU plumbing for the two paths that shard; one process per GPU,
torch.distributed (NCCL over NVLink on the box, gloo in the CPU tests).

* data-parallel mapping: every rank holds a replica of the map and the decoder, takes a contiguous slice
  of the global batch, runs the fused forward/backward kernels on it and exchanges
    - sum  : d loss / d local_geo_features, the packed decoder gradients, the certainty increments
    - max  : local_point_ts_update
  after which every rank applies the identical Adam step (replicated optimiser, no parameter broadcast).
* mesher grid: ranks evaluate disjoint linear ranges of the grid -- no collective (utils/mesher.py here).

The NCCL form of the exchange (`exchange_gradients`) touches no CUDA API, so its logic is testable with gloo on CPU
tensors.  On the GPU box the exchange is `PeerExchange`: gradients accumulate straight into a cudaIpc-shared "peer block"
and ONE kernel per rank (`pinq_peer_exchange`, csrc/peer.cu) does barrier -> owner-sliced reduction over P2P loads
[-> Adam] -> P2P stores -> barrier over NVLink/NVSwitch; NCCL stays as the A/B partner (`PINQ_EXCHANGE=nccl`).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world_info(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_range(n: int, rank: int, world: int):
    """Contiguous slice [first, last) of n items for `rank`; slices tile [0, n) exactly."""
    return n * rank // world, n * (rank + 1) // world


def decimated_indices(first: int, last: int, step: int, device=None):
    """Global indices j in [first, last) with j % step == 0, relative to `first` -- so that the union over ranks of
    shard[rel] equals batch[::step] of the un-sharded batch (reference: utils/mapper.py:484 `coord[::decimation]`)."""
    start = (-first) % step
    if start >= last - first:
        return torch.empty(0, dtype=torch.int64, device=device)
    return torch.arange(start, last - first, step, device=device)


def exchange_gradients(params, sum_buffers=(), max_buffers=(), group=None):
    """All-reduce(sum) of every p.grad (small tensors packed into one flat buffer, large ones in place) and
    of `sum_buffers`; all-reduce(max) of `max_buffers`.  Returns the number of collectives issued."""
    rank, world = world_info(group)
    if world == 1:
        return 0
    calls = 0
    small, large = [], []
    for p in params:
        if p.grad is None:
            continue
        (small if p.grad.numel() <= 65536 else large).append(p.grad)
    for g in large:
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
        calls += 1
    if small:
        flat = torch.cat([g.reshape(-1) for g in small])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        calls += 1
        off = 0
        for g in small:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
    for b in sum_buffers:
        dist.all_reduce(b, op=dist.ReduceOp.SUM, group=group)
        calls += 1
    for b in max_buffers:
        dist.all_reduce(b, op=dist.ReduceOp.MAX, group=group)
        calls += 1
    return calls


def dp_mapping_iteration(mapper, coord, sdf_label, ts, weight, opt, group=None, peer=None, shard=None, keep_grads=None):
    """One data-parallel iteration of Mapper.mapping() (reference: utils/mapper.py:443-573) over a GLOBAL batch
    that every rank holds (indices are drawn with a shared seed): slice -> fused fwd/bwd -> exchange -> Adam.
    The per-rank loss is normalised by the GLOBAL counts so that the summed gradients equal the single-GPU
    gradient of the mean losses.  `peer` (a PeerExchange) selects the peer-memory kernel for the exchange; `shard` =
    (rank, world) overrides the slicing (shard=(0, 1): the un-sharded iteration, no exchange -- self-checks);
    `keep_grads` (a list) receives clones of the exchanged feature and decoder gradients."""
    from .utils.loss import sdf_bce_loss

    cfg = mapper.config
    npts = mapper.neural_points
    rank, world = shard if shard is not None else world_info(group)
    n = coord.shape[0]
    first, last = shard_range(n, rank, world)
    c, lab, t, w = coord[first:last], sdf_label[first:last], ts[first:last], torch.abs(weight[first:last]).detach()
    cert_before = npts.local_point_certainties.clone()
    pred = npts.query_sdf(c, mapper.geo_mlp, t, True, True)["sdf"]
    loss = sdf_bce_loss(pred, lab, mapper.sdf_scale, w, cfg.loss_weight_on, bce_reduction="sum") / n
    if cfg.ekional_loss_on and cfg.weight_e > 0:
        rel = decimated_indices(first, last, cfg.gradient_decimation, device=c.device)
        n_dec_global = (n + cfg.gradient_decimation - 1) // cfg.gradient_decimation
        if rel.numel() > 0:
            g = mapper.get_numerical_gradient(c[rel], pred[rel], cfg.voxel_size_m * cfg.num_grad_step_ratio)
            loss = loss + cfg.weight_e * ((g.norm(2, dim=-1) - 1.0) ** 2).sum() / n_dec_global
    params = [p for grp in opt.param_groups for p in grp["params"]]
    opt.zero_grad(set_to_none=True)
    if peer is not None:  # gradients accumulate in the peer block; one kernel sums them over NVLink (csrc/peer.cu)
        mp = MappingPeer(peer, npts.local_geo_features.shape[0], npts.local_geo_features.shape[1], mapper.geo_mlp)
        mp.attach(npts, mapper.geo_mlp)
        try:
            loss.backward()
        finally:
            MappingPeer.detach(npts, mapper.geo_mlp)
        mp.cert_delta.copy_(npts.local_point_certainties - cert_before)
        mp.ts_update.copy_(npts.local_point_ts_update)
        peer.exchange(n_sum=mp.n_sum, n_max=mp.n_max)
        npts.local_point_certainties = cert_before + mp.cert_delta
        npts.local_point_ts_update.copy_(mp.ts_update)
        fused = list(mapper.geo_mlp.fused_params())
        if keep_grads is not None:
            keep_grads.extend([mp.grad_features.clone()] + [g.clone() for g in mp.grad_decoder])
        for p in params:
            if p is npts.local_geo_features:
                p.grad = mp.grad_features
            else:
                for t, g in zip(fused, mp.grad_decoder):
                    if p is t:
                        p.grad = g
        opt.step()
        opt.zero_grad(set_to_none=True)  # the views die with this iteration's carve
        return loss.detach()
    loss.backward()
    for p in params:  # a rank whose slice never touched a parameter still has to join the collective
        if p.requires_grad and p.grad is None and p.numel() > 0:
            p.grad = torch.zeros_like(p)
    if world > 1:
        delta = npts.local_point_certainties - cert_before
        exchange_gradients(params, sum_buffers=[delta], max_buffers=[npts.local_point_ts_update], group=group)
        npts.local_point_certainties = cert_before + delta
    if keep_grads is not None:
        fused = list(mapper.geo_mlp.fused_params())
        keep_grads.extend([npts.local_geo_features.grad.clone()] + [t.grad.clone() for t in fused])
    opt.step()
    return loss.detach()


# ---------------------------------------------------------------------------------------------------------------
# Peer-memory exchange (include/pinq.h "data-parallel exchange over NVLink peer memory")
def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class _RawCuda:
    """A bare device allocation seen through __cuda_array_interface__ (torch.as_tensor keeps this object alive)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerLayout:
    """Byte layout of a peer block: [pad | fp32 sum region | (fp32 parameter region) | int64 max region]."""

    def __init__(self, n_sum: int, n_max: int = 0, n_param: int = 0):
        from . import _pinq as P
        # every region starts on a 256-byte boundary: the query kernels gather 32-byte feature rows with 256-bit loads
        # straight out of the parameter region
        self.n_sum, self.n_max, self.n_param = _align(n_sum, 64), _align(n_max, 32), _align(n_param, 64)
        self.sum_off = P.PEER_PAD_BYTES
        self.param_off = self.sum_off + 4 * self.n_sum
        self.max_off = self.param_off + 4 * self.n_param
        self.nbytes = _align(self.max_off + 8 * self.n_max, 256)


class PeerExchange:
    """One peer block per rank, mapped into every rank of the node.

    `ranks_on_this_device` > 1 builds that many blocks in ONE process on one GPU and runs their kernels on separate
    streams: the single-GPU form used by the GPU tests (same kernel, same flags, same P2P code path minus the link)."""

    def __init__(self, layout: PeerLayout, device, group=None, simulate_world: int = 0, grid: int = 0, timeout_ms: float = 0.0,
                 cooperative: bool = True, threads: int = 0, multicast=None):
        import ctypes as C
        from . import _pinq as P
        self.P, self.lib, self.layout, self.device = P, P.load(), layout, torch.device(device)
        import os
        self.grid = int(grid)
        # cooperative launch = all CTAs of the exchange resident at once, guaranteed at launch time.  A caller that overlaps the
        # exchange with kernels that do NOT depend on it (the next step's search) passes False: a cooperative grid would wait
        # for a completely free GPU, i.e. behind that search, while a plain launch lets its CTAs take SM slots as they free up
        # (they spin at the flag barrier until the rest arrive, which they do because the other kernel finishes regardless).
        self.cooperative = bool(cooperative)
        # threads per CTA (0 = 512).  An overlapped exchange wants 128: such a CTA fits into the registers one retiring CTA of the
        # neighbour search frees, a 512-thread CTA (100-128 registers per thread) has to wait for an SM that drained completely
        self.threads = int(threads)
        # a rank that does not arrive within this bound makes the kernel give up and set the block's error word
        self.timeout_ms = float(timeout_ms) if timeout_ms else float(os.environ.get("PINQ_PEER_TIMEOUT_MS", "0"))
        self.simulated = simulate_world > 0
        self._mapped, self._owned = [], []
        # NVLS: blocks allocated through torch's symmetric memory (cuMemCreate + one multicast object bound to every rank's
        # block; torch does the handle exchange) instead of cudaMalloc + cudaIpc.  multicast True = required, False = never,
        # None = PINQ_PEER_MULTICAST (0 / 1) or, unset, "where it was measured to win and the box offers it": 8 ranks (the
        # owned slice takes 2 memory instructions per 16 bytes instead of 16; step under the overlapped exchange 1.766 against
        # 1.780 ms) -- at 2 ranks the switch round trip makes it slower (1.756 against 1.721 ms).  self.mc_block == 0 = P2P form.
        self.mc_block, self._symm = 0, None
        if multicast is None:
            env = os.environ.get("PINQ_PEER_MULTICAST", "auto")
            multicast = None if env == "auto" else env not in ("0", "false", "False")
        if not self.simulated and multicast is not False:
            w = world_info(group)[1]
            if w > 1 and (multicast is True or w >= 8):
                self._symm = self._symmetric_block(layout.nbytes, group, required=multicast is True)
        with torch.cuda.device(self.device):
            if self._symm is not None:
                self.rank, self.world = world_info(group)
                tensor, hdl = self._symm
                self.blocks = [int(b) for b in hdl.buffer_ptrs]
                self.mc_block = int(hdl.multicast_ptr)
                if self.blocks[self.rank] != tensor.data_ptr():
                    raise self.P.PinqError("symmetric memory: this rank's buffer is not the tensor it allocated")
                self._epochs = [0]
                self._views = [tensor]
                return
            if self.simulated:
                self.rank, self.world = 0, int(simulate_world)
                self.blocks = []
                for _ in range(self.world):
                    ptr = C.c_void_p()
                    P.check(self.lib.pinq_peer_alloc(layout.nbytes, C.byref(ptr)))
                    self._owned.append(ptr.value)
                    self.blocks.append(ptr.value)
                self.streams = [torch.cuda.Stream(device=self.device) for _ in range(self.world)]
            else:
                self.rank, self.world = world_info(group)
                ptr = C.c_void_p()
                P.check(self.lib.pinq_peer_alloc(layout.nbytes, C.byref(ptr)))
                self._owned.append(ptr.value)
                handle = (C.c_ubyte * P.PEER_HANDLE_BYTES)()
                P.check(self.lib.pinq_peer_export(ptr, handle))
                handles = [None] * self.world
                if self.world > 1:
                    dist.all_gather_object(handles, bytes(handle), group=group)
                self.blocks = []
                for r in range(self.world):
                    if r == self.rank:
                        self.blocks.append(ptr.value)
                        continue
                    buf = (C.c_ubyte * P.PEER_HANDLE_BYTES).from_buffer_copy(handles[r])
                    mapped = C.c_void_p()
                    P.check(self.lib.pinq_peer_open(buf, C.byref(mapped)))
                    self._mapped.append(mapped.value)
                    self.blocks.append(mapped.value)
        self._epochs = [0] * (self.world if self.simulated else 1)
        self._views = [torch.as_tensor(_RawCuda(b, layout.nbytes), device=self.device)
                       for b in (self.blocks if self.simulated else [self.blocks[self.rank]])]

    def _symmetric_block(self, nbytes: int, group, required: bool):
        """(tensor, handle) of a zero-filled symmetric-memory block with an NVLS multicast address, or None when the box has
        no multicast support (every rank takes the same branch: the support is a property of the node)."""
        try:
            import torch.distributed._symmetric_memory as symm
            from torch._C._distributed_c10d import _SymmetricMemory as S
            ok = bool(S.has_multicast_support(torch._C._autograd.DeviceType.CUDA, self.device.index or 0))
        except Exception:  # noqa: BLE001 -- torch build without symmetric memory
            ok = False
        flag = torch.tensor([1 if ok else 0], device=self.device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            if required:
                raise self.P.PinqError("NVLS multicast requested but not supported on this node")
            return None
        t = hdl = None
        try:
            with torch.cuda.device(self.device):
                t = symm.empty(nbytes, dtype=torch.uint8, device=self.device)
                t.zero_()
                torch.cuda.synchronize(self.device)
                hdl = symm.rendezvous(t, group if group is not None else dist.group.WORLD)
            ok = 1 if int(hdl.multicast_ptr) != 0 else 0
        except Exception:  # noqa: BLE001 -- e.g. no permission for the handle exchange inside a sandbox: the P2P form still works
            if required:
                raise
            ok = 0
        flag = torch.tensor([ok], device=self.device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 0:
            if required:
                raise self.P.PinqError("symmetric memory rendezvous returned no multicast address")
            return None
        dist.barrier(group=group)
        return t, hdl

    # -- typed views of a block's regions (this rank's block; `r` selects the block in the simulated form)
    def _view(self, r=None):
        return self._views[r if self.simulated else 0]

    def sum_region(self, r=None):
        L = self.layout
        return self._view(r)[L.sum_off:L.sum_off + 4 * L.n_sum].view(torch.float32)

    def param_region(self, r=None):
        L = self.layout
        return self._view(r)[L.param_off:L.param_off + 4 * L.n_param].view(torch.float32)

    def max_region(self, r=None):
        L = self.layout
        return self._view(r)[L.max_off:L.max_off + 8 * L.n_max].view(torch.int64)

    def next_epochs(self, rank: int, count: int = 1) -> int:
        """First of `count` consecutive epochs for `rank`'s next calls (every rank draws the same sequence)."""
        i = rank if self.simulated else 0
        first = self._epochs[i] + 1
        self._epochs[i] += count
        return first

    def make_args(self, rank, epoch, n_sum=None, n_max=None, sum_off_elems=0, adam=None):
        """pinq_peer_args of one call: the sum region restricted to [sum_off_elems, sum_off_elems + n_sum), the max
        region to its first n_max elements.  adam = dict(exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, step): fused
        optimizer over the first adam['n'] elements (parameters: the head of the param region)."""
        P, L = self.P, self.layout
        n_sum = L.n_sum - sum_off_elems if n_sum is None else _align(int(n_sum), 4)
        n_max = L.n_max if n_max is None else _align(int(n_max), 2)
        if sum_off_elems % 4 or sum_off_elems + n_sum > L.n_sum or n_max > L.n_max:
            raise P.PinqError("exchange outside the peer block's regions")
        a = P.PeerArgs()
        a.rank, a.world = rank, self.world
        for r, b in enumerate(self.blocks):
            a.block[r] = b
        a.mc_block = self.mc_block or None
        a.sum_off, a.n_sum, a.max_off, a.n_max = L.sum_off + 4 * sum_off_elems, n_sum, L.max_off, n_max
        a.epoch, a.grid, a.timeout_ms, a.threads = epoch, self.grid, self.timeout_ms, self.threads
        a.cooperative = 0 if (self.simulated or not self.cooperative) else 1  # simulated ranks share one device and wait for each other
        if adam is not None:
            adam = dict(adam)
            adam.setdefault("n", L.n_param)
            if sum_off_elems or _align(int(adam["n"]), 4) > min(n_sum, L.n_param):
                raise P.PinqError("fused Adam works on the head of the sum / param regions")
            a.param_off, a.n_adam = L.param_off, _align(int(adam["n"]), 4)
            a.exp_avg, a.exp_avg_sq = P.ptr(adam["exp_avg"]), P.ptr(adam["exp_avg_sq"])
            a.lr, a.beta1, a.beta2, a.adam_eps, a.adam_step = adam["lr"], adam["beta1"], adam["beta2"], adam["eps"], adam["step"]
        return a

    def exchange(self, adam=None, n_sum=None, n_max=None, sum_off_elems=0):
        """Sum / max the regions over the ranks (and, with `adam`, apply the optimizer step on the owned slice and
        broadcast the parameters).  `n_sum` / `n_max` / `sum_off_elems` restrict the call to a part of the regions
        (identical on every rank).  Asynchronous on the current stream; every rank calls it in lock-step.
        Simulated form: `adam['exp_avg'] / ['exp_avg_sq']` are lists with one tensor per simulated rank."""
        P = self.P
        if not self.simulated:
            a = self.make_args(self.rank, self.next_epochs(self.rank), n_sum, n_max, sum_off_elems, adam)
            P.check(self.lib.pinq_peer_exchange(a, P.stream()))
            return
        cur = torch.cuda.current_stream(self.device)
        for r, st in enumerate(self.streams):  # the "ranks" must be co-resident: small grids on separate streams
            st.wait_stream(cur)
            ad = None if adam is None else dict(adam, exp_avg=adam["exp_avg"][r], exp_avg_sq=adam["exp_avg_sq"][r])
            a = self.make_args(r, self.next_epochs(r), n_sum, n_max, sum_off_elems, ad)
            P.check(self.lib.pinq_peer_exchange(a, st.cuda_stream))
        for st in self.streams:
            cur.wait_stream(st)

    def error(self) -> int:
        """Non-zero if a barrier timed out in any exchange so far (synchronises)."""
        off = int(self.lib.pinq_peer_error_offset())
        return int(max(int(v[off:off + 4].view(torch.int32).item()) & 0x7fffffff for v in self._views))

    def close(self):
        self._views = []
        self._symm = None
        for m in self._mapped:
            self.lib.pinq_peer_close(m)
        for o in self._owned:
            self.lib.pinq_peer_free(o)
        self._mapped, self._owned = [], []


class MappingPeer:
    """Carves the sum / max regions of a PeerExchange for one data-parallel mapping iteration:
    sum = [d loss / d local_geo_features (M+1)*F | dW1 | db1 | dW2 | db2 | pad | certainty increments M], max = [ts_update M]."""

    def __init__(self, peer: PeerExchange, feature_rows: int, feature_dim: int, decoder, r=None):
        self.peer = peer
        sumr, maxr = peer.sum_region(r), peer.max_region(r)
        self._sum = sumr
        nf = feature_rows * feature_dim
        self.grad_features = sumr[:nf].view(feature_rows, feature_dim)
        off, self.grad_decoder = nf, []
        for t in (decoder.fused_params() if decoder is not None else ()):  # None: frozen decoder (H11), features only
            self.grad_decoder.append(sumr[off:off + t.numel()].view(t.shape))
            off += t.numel()
        self.n_grad = off                      # feature + decoder gradients: the fused optimizer's element count
        off = _align(off, 4)
        self.cert_off = off
        m = feature_rows - 1
        self.cert_delta = sumr[off:off + m]
        self.n_sum = off + m
        self.ts_update = maxr[:m]
        self.n_max = m
        if self.n_sum > peer.layout.n_sum or m > peer.layout.n_max:
            raise peer.P.PinqError("peer block too small for this local map")

    @staticmethod
    def capacity(feature_rows: int, feature_dim: int, decoder, with_params: bool = False):
        """Layout for local maps of up to `feature_rows` rows; `with_params`: also a parameter region (fused Adam)."""
        nd = sum(t.numel() for t in decoder.fused_params()) if decoder is not None else 0
        n_grad = _align(feature_rows * feature_dim + nd, 4)
        return PeerLayout(n_grad + feature_rows + 8, feature_rows + 2, n_grad if with_params else 0)

    def attach(self, npts, decoder):
        """Zeroes the used prefix and points the fused backward's gradient sinks at it."""
        self._sum[:self.n_sum].zero_()
        npts.feature_grad_sink = self.grad_features
        decoder.grad_sink = tuple(self.grad_decoder)

    @staticmethod
    def detach(npts, decoder):
        npts.feature_grad_sink = None
        decoder.grad_sink = None


class DataParallelMapping:
    """Data-parallel form of Mapper.mapping's device-side loop (`Mapper.dp = DataParallelMapping(peer)`): every rank draws
    the same global batches (shared seed), trains on its slice, and `pinq_mapping_run` replaces its Adam launch by
    `pinq_peer_exchange` with the fused optimizer -- per iteration ONE extra kernel sums the gradient blocks over NVLink,
    steps the owned slice and broadcasts the parameters.  Certainty increments and ts_update, which no iteration reads,
    are exchanged once after the last iteration.  `rank` is only given in the simulated single-GPU form."""

    def __init__(self, peer: PeerExchange, rank=None):
        self.peer = peer
        self.rank = peer.rank if rank is None else int(rank)
        self.world = peer.world
        self._keep = None

    def shard(self, n: int):
        return shard_range(n, self.rank, self.world)

    def begin(self, npts, decoder):
        """Moves the trained parameters into this rank's peer block (param region) and carves the gradient block."""
        peer = self.peer
        r = self.rank if peer.simulated else None
        rows, fdim = npts.local_geo_features.shape
        carve = MappingPeer(peer, rows, fdim, decoder, r=r)
        par = peer.param_region(r)
        if _align(carve.n_grad, 4) > par.numel():
            raise peer.P.PinqError("peer block has no parameter region of this size (MappingPeer.capacity(with_params=True))")
        tensors = [npts.local_geo_features] + (list(decoder.fused_params()) if decoder is not None else [])
        home, off = [], 0
        for t in tensors:
            cnt = t.numel()
            view = par[off:off + cnt].view(t.shape)
            view.copy_(t.data)
            home.append(t.data)
            t.data = view
            off += cnt
        carve._sum[:carve.n_sum].zero_()
        return {"carve": carve, "grads": carve._sum, "tensors": tensors, "home": home,
                "cert_before": npts.local_point_certainties.clone()}

    def peer_args(self, iters: int, n_grad: int, exp_avg, exp_avg_sq):
        adam = {"exp_avg": exp_avg, "exp_avg_sq": exp_avg_sq, "n": n_grad, "lr": 0.0, "beta1": 0.0, "beta2": 0.0, "eps": 0.0,
                "step": 1}  # hyper-parameters and step are filled in per iteration by pinq_mapping_run
        self._keep = self.peer.make_args(self.rank, self.peer.next_epochs(self.rank, iters), n_sum=_align(n_grad, 4), n_max=0,
                                         adam=adam)
        return self._keep

    def end(self, npts, decoder, moved):
        peer, carve = self.peer, moved["carve"]
        m = carve.n_max
        carve.cert_delta.copy_(npts.local_point_certainties - moved["cert_before"])
        carve.ts_update.copy_(npts.local_point_ts_update)
        a = peer.make_args(self.rank, peer.next_epochs(self.rank), n_sum=m, n_max=m, sum_off_elems=carve.cert_off)
        peer.P.check(peer.lib.pinq_peer_exchange(a, peer.P.stream()))
        npts.local_point_certainties = moved["cert_before"] + carve.cert_delta
        npts.local_point_ts_update.copy_(carve.ts_update)
        self.abort(moved)
        # a rank that never arrived at a barrier leaves partial sums / parameters behind: fail loudly (the parameters are
        # back in torch-owned memory, so the caller can still inspect or checkpoint them)
        # (simulated ranks of the tests are launched back to back from one host thread and meet on the device: no host
        # synchronisation may sit between them, the test reads the error word itself)
        err = 0 if peer.simulated else peer.error()
        if err:
            raise peer.P.PinqError("pinq_peer_exchange: barrier timeout in epoch %d of this mapping call -- a rank was late or "
                                   "dead; the trained parameters of this call are not the data-parallel result" % (err >> 1))

    @staticmethod
    def abort(moved):
        """Parameters (with whatever values they hold now) back into the torch-owned memory they came from."""
        for t, h in zip(moved["tensors"], moved["home"]):
            if t.data.data_ptr() != h.data_ptr():
                h.copy_(t.data)
                t.data = h
