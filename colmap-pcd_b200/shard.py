"""Host-side sharding of the association path over one process per GPU.

The cloud and its index are replicated; SfM points (queries) are split into contiguous shards, images
round-robin.  The path has exactly one exchange step per pass: the per-match [r, Jx, Jy, Jz] blocks are
gathered and sum r^2 / the accepted and ground counters are all-reduced.  Everything here works on any
torch.distributed backend (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def shard_bounds(n, world, rank):
    """Contiguous, balanced split of n units: the first n % world shards get one extra."""
    base, extra = divmod(int(n), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_images(n_img, world, rank):
    """Round-robin image assignment for the depth projection (450 images -> 57/56 per GPU at world 8)."""
    return list(range(rank, int(n_img), int(world)))


def reduce_stats(sum_sq, accepted, ground, group=None):
    """All-reduce (sum, fp64) of the residual norm and the two counters.  Returns a float64 tensor [3]."""
    t = sum_sq if torch.is_tensor(sum_sq) and sum_sq.numel() == 3 else None
    if t is None:
        dev = sum_sq.device if torch.is_tensor(sum_sq) else "cpu"
        t = torch.tensor([float(sum_sq), float(accepted), float(ground)], dtype=torch.float64, device=dev)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


class Exchange:
    """The path's one exchange step per pass: all-gather of the per-rank residuals r [Q] and Jacobian blocks J [Q, 3]
    (equal Q on every rank) and all-reduce (sum) of the fp64 statistics [sum r^2, accepted, ground].

    On CUDA everything runs on a side stream, so that the collectives of pass k overlap the kernels of pass k+1;
    the caller alternates between two (r, J) output buffers and calls wait_buffer(b) before overwriting buffer b.
    """

    def __init__(self, q_per_rank, device, group=None, root=0, wire="f64", pack=None):
        """root = r: gather to rank r only (the consumer is the host-side solver of one process; the other ranks
        send their blocks once and receive nothing).  root = None: all-gather, every rank gets every block.
        wire = "f64": r and J travel as computed (32 B/query).  wire = "f32": `pack(r, J, out, stream_ptr)` (the
        library's pcd_pack_blocks_f32_dev) rounds each block to a float4 {r, J0, J1, J2} first -- 16 B/query on the
        wire, relative error <= 2^-24; the gathered views are float32 and the fp64 blocks stay with their producer."""
        if wire not in ("f64", "f32") or (wire == "f32" and pack is None):
            raise ValueError("Exchange: wire is 'f64' or 'f32' (the latter needs the pack callable)")
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group = group
        self.root = root
        self.wire, self.pack = wire, pack
        self.q = int(q_per_rank)
        dev = torch.device(device)
        self.cuda = dev.type == "cuda"
        full = self.world * self.q if (root is None or self.rank == root) else 0
        if wire == "f32":
            self.w_local = [torch.empty((self.q, 4), dtype=torch.float32, device=dev) for _ in range(2)]
            self.w_all = torch.empty((full, 4), dtype=torch.float32, device=dev)
            self.r_all, self.J_all = self.w_all[:, 0], self.w_all[:, 1:4]
        else:
            self.r_all = torch.empty(full, dtype=torch.float64, device=dev)
            self.J_all = torch.empty((full, 3), dtype=torch.float64, device=dev)
        self.stats = torch.zeros(3, dtype=torch.float64, device=dev)
        # high priority: the collective's CTAs must not queue behind the search kernels' 60 k-block grids
        self.stream = torch.cuda.Stream(device=dev, priority=-1) if self.cuda else None
        self.events = [torch.cuda.Event() for _ in range(2)] if self.cuda else None
        self.pending = [False, False]

    def _run(self, r, J, stats, buf=0):
        parts = [(self.r_all, r), (self.J_all, J)]
        if self.wire == "f32":
            w = self.w_local[buf]
            self.pack(r, J, w, self.stream.cuda_stream if self.cuda else 0)
            parts = [(self.w_all, w)]
        if self.world > 1:
            for dst, src in parts:
                if self.root is None:
                    dist.all_gather_into_tensor(dst, src, group=self.group)
                else:
                    here = self.rank == self.root
                    dist.gather(src, list(dst.split(self.q)) if here else None, dst=self.root, group=self.group)
            self.stats.copy_(stats)
            dist.all_reduce(self.stats, op=dist.ReduceOp.SUM, group=self.group)
        else:
            for dst, src in parts:
                dst.copy_(src)
            self.stats.copy_(stats)

    def start(self, r, J, stats, buf=0):
        """r, J: this rank's outputs of the pass just enqueued; stats: fp64 [3]."""
        if not self.cuda:
            self._run(r, J, stats, buf)
            return
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            self._run(r, J, stats, buf)
            self.events[buf].record(self.stream)
        self.pending[buf] = True

    def wait_buffer(self, buf):
        """Make the current stream wait until the exchange that read buffer `buf` has finished."""
        if self.cuda and self.pending[buf]:
            torch.cuda.current_stream().wait_event(self.events[buf])
            self.pending[buf] = False

    def wait(self):
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.stream)
        return self.r_all, self.J_all, self.stats


class PeerExchange:
    """The same exchange step written over NVLink peer memory instead of NCCL's all-gather kernels.

    Every rank owns a symmetric receive buffer (torch symmetric memory: each peer's buffer is mapped into the local
    address space).  After a pass, a rank pushes its r [Q] and J [Q, 3] blocks straight into its slot of every peer's
    buffer with device-to-device copies on per-peer side streams -- copy engines over NVLink / NVSwitch, no SM taken
    from the search kernels that run the next pass at the same time -- then joins a signal-pad barrier, after which
    every rank holds every block.  Two receive buffers alternate.  Ordering between passes: the pushes of pass k + 2
    into receive buffer b wait for the event recorded behind the barrier of pass k + 1, i.e. until EVERY rank has
    entered start(k + 1); a consumer must therefore enqueue its reads of pass k (wait_buffer / wait, then its kernels)
    BEFORE it calls start(k + 1) -- after that call the buffer of pass k may be overwritten by a faster peer.
    The fp64 statistics still go through a (tiny) NCCL all-reduce.
    wire = "f32": the blocks are first rounded to float4 {r, J0, J1, J2} by `pack` (see Exchange) and one 16 B/query
    copy per peer replaces the two fp64 ones.
    """

    def __init__(self, q_per_rank, device, group=None, wire="f64", pack=None, chunks=1):
        """chunks > 1: every peer's copy is cut into that many pieces issued on streams of their own (more copies in
        flight per peer)."""
        import torch.distributed._symmetric_memory as symm
        if wire not in ("f64", "f32") or (wire == "f32" and pack is None):
            raise ValueError("PeerExchange: wire is 'f64' or 'f32' (the latter needs the pack callable)")
        self.wire, self.pack = wire, pack
        self.chunks = max(1, int(chunks))
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.q = int(q_per_rank)
        dev = torch.device(device)
        n = self.world * self.q
        self.recv, self.hdl, self.peer_views = [], [], []
        wdt = torch.float32 if wire == "f32" else torch.float64
        for _ in range(2):
            t = symm.empty(4 * n, dtype=wdt, device=dev)
            h = symm.rendezvous(t, group=self.group)
            self.recv.append(t); self.hdl.append(h)
            self.peer_views.append([h.get_buffer(p, (4 * n,), wdt) for p in range(self.world)])
        if wire == "f32":
            self.w_local = [torch.empty((self.q, 4), dtype=torch.float32, device=dev) for _ in range(2)]
            self.pack_stream = torch.cuda.Stream(device=dev, priority=-1)
        self.stats = torch.zeros(3, dtype=torch.float64, device=dev)
        self.main = torch.cuda.Stream(device=dev, priority=-1)
        self.streams = [torch.cuda.Stream(device=dev, priority=-1) for _ in range(self.world * self.chunks)]
        self.events = [torch.cuda.Event() for _ in range(2)]
        self.pending = [False, False]
        self.recorded = [False, False]
        self.last = 0

    def views(self, buf):
        n = self.world * self.q
        if self.wire == "f32":
            w = self.recv[buf].view(n, 4)
            return w[:, 0], w[:, 1:4]
        return self.recv[buf][:n], self.recv[buf][n:].view(n, 3)

    def start(self, r, J, stats, buf=0):
        n, q, me = self.world * self.q, self.q, self.rank
        cur = torch.cuda.current_stream()
        src = cur
        if self.wire == "f32":
            self.pack_stream.wait_stream(cur)
            with torch.cuda.stream(self.pack_stream):
                self.pack(r, J, self.w_local[buf], self.pack_stream.cuda_stream)
            src, wf = self.pack_stream, self.w_local[buf].reshape(-1)
        Jf = J.reshape(-1)
        C = self.chunks
        for k in range(self.world):                      # own slot first, then the peers in ring order
            p = (me + k) % self.world
            for c in range(C):
                a, b = q * c // C, q * (c + 1) // C      # this piece's rows
                st = self.streams[p * C + c]
                st.wait_stream(src)
                if self.recorded[1 - buf]:
                    st.wait_event(self.events[1 - buf])  # every rank is past the barrier of the previous pass
                with torch.cuda.stream(st):
                    dst = self.peer_views[buf][p]
                    if self.wire == "f32":
                        dst[4 * (me * q + a):4 * (me * q + b)].copy_(wf[4 * a:4 * b], non_blocking=True)
                    else:
                        dst[me * q + a:me * q + b].copy_(r[a:b], non_blocking=True)
                        dst[n + 3 * (me * q + a):n + 3 * (me * q + b)].copy_(Jf[3 * a:3 * b], non_blocking=True)
        self.main.wait_stream(cur)
        for st in self.streams:
            self.main.wait_stream(st)
        with torch.cuda.stream(self.main):
            self.hdl[buf].barrier(channel=0)             # every rank's pushes of this pass have landed
            self.stats.copy_(stats)
            dist.all_reduce(self.stats, op=dist.ReduceOp.SUM, group=self.group)
            self.events[buf].record(self.main)
        self.pending[buf] = True
        self.recorded[buf] = True
        self.last = buf

    def wait_buffer(self, buf):
        if self.pending[buf]:
            torch.cuda.current_stream().wait_event(self.events[buf])
            self.pending[buf] = False

    def wait(self):
        torch.cuda.current_stream().wait_stream(self.main)
        r_all, J_all = self.views(self.last)
        return r_all, J_all, self.stats


class HybridExchange:
    """Rows [0, split) of every rank's blocks travel by NCCL all-gather (its CTAs capped by NCCL_MAX_CTAS so that the
    search kernels keep their SMs), rows [split, Q) by copy-engine pushes into symmetric memory: the two paths use
    different hardware (SMs + NVLink vs copy engines + NVLink) and run side by side.  fp64 blocks on both."""

    def __init__(self, q_per_rank, device, group=None, nccl_frac=0.45, chunks=1):
        self.q = int(q_per_rank)
        self.split = max(1, min(self.q - 1, int(self.q * float(nccl_frac))))
        self.a = Exchange(self.split, device, group=group, root=None)
        self.b = PeerExchange(self.q - self.split, device, group=group, chunks=chunks)
        self.world = self.a.world

    def start(self, r, J, stats, buf=0):
        s = self.split
        self.a.start(r[:s], J[:s], stats, buf)
        self.b.start(r[s:], J[s:], stats, buf)

    def wait_buffer(self, buf):
        self.a.wait_buffer(buf); self.b.wait_buffer(buf)

    def wait(self):
        ra, Ja, st = self.a.wait()
        rb, Jb, _ = self.b.wait()
        s, q, w = self.split, self.q, self.world
        r_all = torch.cat([torch.cat([ra[k * s:(k + 1) * s], rb[k * (q - s):(k + 1) * (q - s)]]) for k in range(w)])
        J_all = torch.cat([torch.cat([Ja[k * s:(k + 1) * s], Jb[k * (q - s):(k + 1) * (q - s)]]) for k in range(w)])
        return r_all, J_all, st


def gather_ragged(local, group=None):
    """All-gather of per-rank 1-D/2-D tensors whose first dimension differs (result sets of a ragged shard)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros((m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[:s] for p, s in zip(parts, sizes)], 0)
