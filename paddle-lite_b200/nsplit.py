"""Multi-GPU host logic.

* conv nets: `shard_batch` -- images are independent units, weights replicated, no
  collective on the per-image path (one Predictor per thread in the reference,
  lite/api/lite_multithread_test.cc).
* large fc / GEMM: `NSplitGemm` -- column-parallel: rank r owns B[:, n0:n1] with its
  scale/bias slice and a full copy of A, computes C_r[M, n1-n0] with the tcgen05 GEMM
  and all-gathers the column blocks (NCCL over NVLink; gloo in the CPU tests).

torch / torch.distributed are plumbing only; the GEMM goes through the C-ABI.
"""
import torch
import torch.distributed as dist


def shard_batch(total, world, rank):
    """[begin, end) of the images rank owns; remainders go to the low ranks."""
    if world <= 0 or not (0 <= rank < world) or total < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(total, world)
    b = rank * base + min(rank, rem)
    return b, b + base + (1 if rank < rem else 0)


def shard_columns(n, world, rank, align=16):
    """[n0, n1) of the output columns rank owns.  Every block but the last is the same
    multiple of `align` wide (16 = one TMA row of the K-major weight slice and one
    16-byte store), so the gathered layout is regular."""
    if world <= 0 or not (0 <= rank < world) or n <= 0:
        raise ValueError("bad shard request")
    per = -(-n // world)
    per = -(-per // align) * align
    n0 = min(rank * per, n)
    return n0, min(n0 + per, n)


def block_width(n, world, align=16):
    per = -(-n // world)
    return -(-per // align) * align


def gather_columns(local_c, n, group=None):
    """local_c [M, per] (this rank's column block, padded to the common block width)
    -> C [M, n] on every rank.  One all_gather_into_tensor ([G, M, per], block layout)
    followed by one de-interleaving copy."""
    world = dist.get_world_size(group)
    M, per = local_c.shape
    blocks = torch.empty((world * M, per), dtype=local_c.dtype, device=local_c.device)
    dist.all_gather_into_tensor(blocks, local_c.contiguous(), group=group)
    return blocks.view(world, M, per).permute(1, 0, 2).reshape(M, world * per)[:, :n].contiguous()


class NSplitGemm:
    """C[M,N] = q(scale[n] * (A[M,K] . B[K,N]) + bias[n]) with B split by columns."""

    def __init__(self, pl, B_kn, scale, bias, out_dtype, act=0, act_arg=0.0, group=None, device=None):
        self.pl = pl
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        K, N = B_kn.shape
        self.K, self.N = K, N
        self.per = block_width(N, self.world)
        self.n0, self.n1 = shard_columns(N, self.world, self.rank)
        self.out_dtype = out_dtype
        self.ldk = -(-K // 16) * 16
        nb = max(self.n1 - self.n0, 0)
        # K-major slice [per, ldk]; rows past n1 stay zero so every rank gathers `per` columns
        bt = torch.zeros((self.per, self.ldk), dtype=torch.int8)
        bt[:nb, :K] = torch.as_tensor(B_kn[:, self.n0:self.n1]).t()
        self.bt = bt.to(self.device)
        sc = torch.zeros(self.per, dtype=torch.float32)
        bi = torch.zeros(self.per, dtype=torch.float32)
        sc[:nb] = torch.as_tensor(scale[self.n0:self.n1], dtype=torch.float32)
        if bias is not None:
            bi[:nb] = torch.as_tensor(bias[self.n0:self.n1], dtype=torch.float32)
        self.scale, self.bias = sc.to(self.device), bi.to(self.device)
        self.act, self.act_arg = act, act_arg
        self._local = None

    def local(self, A_dev):
        """this rank's column block [M, per] (enqueue only on the current stream)."""
        M = A_dev.shape[0]
        assert A_dev.shape[1] == self.ldk and A_dev.dtype == torch.int8
        tdt = torch.int8 if self.out_dtype == self.pl.OUT_INT8 else torch.float32
        if self._local is None or self._local.shape[0] != M:
            self._local = torch.empty((M, self.per), dtype=tdt, device=self.device)
        self.pl.gemm(A_dev, self.bt, self.scale, self.bias, self.act, self.act_arg, self.out_dtype,
                     y=self._local, M=M, N=self.per, K=self.K)
        return self._local

    def __call__(self, A_dev):
        c = self.local(A_dev)
        if self.world == 1:
            return c[:, :self.N]
        return gather_columns(c, self.N, self.group)

    # -- fused GEMM + all-gather over NVLink peer memory (int8 out) --------------------------------
    def fused_setup(self, M):
        """Allocates the full result [M, world * per] in symmetric memory (every rank maps every other
        rank's buffer) -- TWO of them, used alternately -- and uploads the peer tensor maps of both once
        (pl_gemm_i8_peers_prepare), so that a fused() call is one kernel launch + one device-side barrier.

        Why two buffers: rank A's call n+1 stores into rank B's result while B may still be reading the
        result of call n (nothing orders A's peer stores after B's consumer).  With alternating buffers the
        stores of call n+2 are the first to touch buffer (n mod 2) again, and A cannot start call n+2 before
        it has passed the barrier of call n+1, which B only enters after its consumer of call n (stream
        order on B).  No extra barrier is needed."""
        import torch.distributed._symmetric_memory as symm_mem
        int8 = self.out_dtype == self.pl.OUT_INT8  # (fp32 results: the multicast variant only)
        tdt = torch.int8 if int8 else torch.float32
        esz = 1 if int8 else 4
        grp = self.group if self.group is not None else dist.group.WORLD
        off = self.rank * self.per  # this rank's column block inside everybody's result
        self._fulls, self._hdls, self._maps, self._ys = [], [], [], []
        self._n_peers = self.world - 1
        for _ in range(2):
            full = symm_mem.empty((M, self.world * self.per), dtype=tdt, device=self.device)
            hdl = symm_mem.rendezvous(full, grp)
            peer_ptrs = [int(hdl.buffer_ptrs[r]) + off * esz for r in range(self.world) if r != self.rank]
            maps = torch.empty(128 * 16, dtype=torch.uint8, device=self.device)
            y = full[:, off:off + self.per]
            if peer_ptrs and int8:
                self.pl.gemm_peers_prepare(y, peer_ptrs, maps, self.scale, self.bias, M, self.per, self.K,
                                           self.act, self.act_arg)
            self._fulls.append(full)
            self._hdls.append(hdl)
            self._maps.append(maps)
            self._ys.append(y)
        self._fM, self._turn = M, 0
        self._full = self._fulls[0]
        return self._fulls[0]

    def mcast_supported(self):
        """NVSwitch multicast (NVLS) is available for this group's symmetric memory"""
        if getattr(self, "_fM", None) is None:
            return False
        return all(int(h.multicast_ptr) != 0 for h in self._hdls)

    def fused_mcast(self, A_dev):
        """Like fused(), but every staged tile leaves the GPU ONCE, through the multicast address of the result
        buffer (multimem.st in the GEMM's store warp): the switch replicates it into all ranks' results, own
        included.  NVLink egress per GPU is the size of its column block, not (G - 1) times that."""
        M = A_dev.shape[0]
        if getattr(self, "_fM", None) != M:
            self.fused_setup(M)
        t = self._turn
        self._turn ^= 1
        mc = int(self._hdls[t].multicast_ptr)
        if mc == 0:
            raise RuntimeError("this group's symmetric memory has no multicast mapping (no NVLS)")
        if self.out_dtype == self.pl.OUT_INT8:
            self.pl.gemm_mcast(A_dev, self.bt, self.scale, self.bias, self._ys[t], mc + self.rank * self.per,
                               self.act, self.act_arg, N=self.per, K=self.K)
        else:
            # fp32: plain GEMM into this rank's block of its own result, then one coalesced multicast broadcast
            # of that block (multimem.st from the fp32 epilogue's registers is 10x slower: 16-byte pieces of 32
            # different rows per warp store)
            y = self._ys[t]
            self.pl.gemm(A_dev, self.bt, self.scale, self.bias, self.act, self.act_arg, self.out_dtype, y=y, M=M,
                         N=self.per, K=self.K)
            self.pl.mcast_copy_2d(y, mc + self.rank * self.per * 4, self.world * self.per * 4)
        self._hdls[t].barrier()
        return self._fulls[t][:, :self.N]

    def fused(self, A_dev):
        """C[M, N] on every rank: each rank's GEMM kernel TMA-stores its output tiles into its own result
        buffer and straight into the peers' (NVLink), overlapped with the mainloop; one device-side barrier
        of the ranks afterwards.  No collective call, no de-interleaving copy.  The returned view is valid
        until the second next fused() call (two result buffers alternate, see fused_setup)."""
        assert self.out_dtype == self.pl.OUT_INT8, "peer stores exist on the int8 TMA-store epilogue only"
        M = A_dev.shape[0]
        if getattr(self, "_fM", None) != M:
            self.fused_setup(M)
        t = self._turn
        self._turn ^= 1
        self.pl.gemm_peers(A_dev, self.bt, self.scale, self.bias, self._ys[t], self._n_peers, self._maps[t],
                           self.act, self.act_arg, N=self.per, K=self.K)
        self._hdls[t].barrier()
        return self._fulls[t][:, :self.N]
