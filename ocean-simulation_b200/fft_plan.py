"""Executable model of the in-kernel FFT plan used by csrc/ocean_fft.cuh.

The CUDA kernels transform a line of N = 2^m complex points with T = N/8 threads, each thread
holding 8 points in registers.  Stages are radix-8 decimation-in-frequency butterflies done in
registers; between stages the 8 results are exchanged through shared memory (128-bit elements:
two packed spectra per element).  The last stage has radix N / 8^(S-1) in {2, 4, 8}.  Input and
output are both in natural order with thread p holding points p + T*j (j = 0..7), so global
loads and stores are coalesced and no bit-reversal pass exists.

This module restates the index algebra in numpy, one vectorised "thread" axis, so that
  * tests/test_fft_plan.py can prove the plan equals N * numpy.fft.ifft for every supported N
    (the reference network IFFT.compute:48-69 is an unnormalised inverse DFT), and
  * the shared-memory swizzle can be checked for bank conflicts per quarter-warp (a 128-bit
    access is conflict-free when the 8 lanes of a quarter-warp hit 8 distinct 16-byte bank groups).
It is host-side design verification only: nothing here runs on the hot path.
"""
import numpy as np


def stage_radices(N):
    """[8, 8, ..., r_last] with product N and r_last in {2, 4, 8}."""
    assert N >= 16 and N & (N - 1) == 0
    r = []
    rem = N
    while rem > 8:
        r.append(8)
        rem //= 8
    r.append(rem)
    if r[-1] == 1:
        r.pop()
    return r


def swizzle(F, exchange=1, dense0=False):
    """Physical element index (units of one 128-bit element) for logical index F.
    A thread's exchange-0 slots (tid + T*q) are exactly the slots of its own 8 input points, so the
    buffer the inputs are staged in (cp.async, TMA) doubles as the exchange buffer with no barrier.
    dense0: exchange 0 is unswizzled -- needed when a TMA tensor copy lands the tile in natural order
    (column kernel); per-thread cp.async (row kernel) can stage straight into swizzled slots."""
    if exchange == 0 and dense0:
        return F
    return F ^ ((F >> 3) & 7)


def _butterfly(v, r):
    """Inverse DFT of size r along axis 0: u[q] = sum_j v[j] e^{+2 pi i j q / r}."""
    j = np.arange(r)
    return np.exp(2j * np.pi * np.outer(j, j) / r) @ v


def run_plan(x, W=1, check_conflicts=False, dense0=False):
    """Transform W interleaved lines.  x: complex array [W][N].  Returns (X[W][N], conflicts).

    Lane numbering inside a CTA is lane = xw + W * tid (column index fastest), which is what
    makes the column kernel's global accesses contiguous across W adjacent columns.
    Shared-memory logical index for element F of column xw is G = F * W + xw.
    """
    x = np.asarray(x, np.complex128)
    W_, N = x.shape
    assert W_ == W
    T = N // 8
    radices = stage_radices(N)
    S = len(radices)
    tw = np.exp(2j * np.pi * np.arange(N) / N)  # the table W_N[i] the kernel reads
    conflicts = []

    tid = np.arange(T)
    # registers: reg[j][tid][xw]
    reg = np.stack([x[:, tid + T * j].T for j in range(8)])  # stage 0 load: point tid + T*j

    def record(kind, s, G):  # G: [tid][xw] logical indices accessed by one instruction
        if not check_conflicts:
            return
        lane = (np.arange(W)[None, :] + W * tid[:, None]).ravel()
        phys = swizzle(G.ravel(), s, dense0)
        order = np.argsort(lane)
        phys = phys[order]
        worst = 1
        for q in range(0, phys.size, 8):
            grp = np.unique(phys[q:q + 8])
            banks = grp & 7
            worst = max(worst, np.bincount(banks, minlength=8).max())
        conflicts.append((kind, s, int(worst)))

    xw = np.arange(W)[None, :]
    for s in range(S):
        r = radices[s]
        Cs = 8 ** s
        Ms = N // Cs
        last = s == S - 1
        if not last or r == 8:
            # thread -> (c, t): c slow, t fast (for the last radix-8 stage Ms/8 == 1, tid == c)
            c, t = tid // (Ms // 8), tid % (Ms // 8)
            u = np.stack([_butterfly(reg[:, i, :], 8) for i in range(T)], axis=1)  # [q][tid][xw]
            if Ms > 8:
                for q in range(8):
                    u[q] *= tw[(t * q * Cs) % N][:, None]
            if last:
                reg = u  # output X[c + Cs*q] = X[p + T*q]
                break
            # exchange: write A_{s+1}[c + Cs*q][t]
            Mn = Ms // 8
            smem = np.zeros(N * W, np.complex128)
            for q in range(8):
                F = (c + Cs * q) * Mn + t
                G = F[:, None] * W + xw
                record("write", s, G)
                smem[swizzle(G, s, dense0)] = u[q]
            # read for stage s+1
            rn = radices[s + 1]
            Cn = Cs * 8
            if s + 1 < S - 1 or rn == 8:
                cn, tn = tid // (Mn // 8), tid % (Mn // 8)
                new = []
                for j in range(8):
                    F = cn * Mn + tn + (Mn // 8) * j
                    G = F[:, None] * W + xw
                    record("read", s, G)
                    new.append(smem[swizzle(G, s, dense0)])
                reg = np.stack(new)
            else:
                # last stage with radix rn < 8: thread p handles sub-problems c_i = p + T*i
                new = []
                for i in range(8 // rn):
                    for j in range(rn):
                        F = (tid + T * i) * rn + j  # Mn == rn
                        G = F[:, None] * W + xw
                        record("read", s, G)
                        new.append(smem[swizzle(G, s, dense0)])
                reg = np.stack(new)  # slot i*rn + j
        else:
            # last stage, radix r < 8: 8/r independent butterflies per thread, no twiddle
            out = np.zeros_like(reg)
            per = 8 // r
            for i in range(per):
                v = reg[i * r:(i + 1) * r]
                u = np.stack([_butterfly(v[:, k, :], r) for k in range(T)], axis=1)
                for q in range(r):
                    out[i + per * q] = u[q]  # x = p + T*(i + (8/r) q)
            reg = out
    X = np.zeros((W, N), np.complex128)
    for j in range(8):
        X[:, tid + T * j] = reg[j].T
    return X, conflicts


# ---------------------------------------------------------------------------------------------------------------------
# Shared-memory wavefronts of one column tile, instruction by instruction (ocean_cols_tma_kernel, compute warps).
# A wavefront is one 128-byte pass of the shared-memory pipe; a warp-wide access needs as many passes as the busiest
# bank has distinct 4-byte words (identical addresses are broadcast).  `ideal` is what ncu's source page calls
# "L1 Wavefronts Shared Ideal": ceil(distinct bytes / 128).  tests/test_fft_plan.py compares the totals with the
# per-instruction counters of the committed ncu capture (profiles/r2_ncu_smem_by_instruction.txt).

def warp_wavefronts(byte_addresses, access_bytes):
    """(wavefronts, ideal) of one warp-wide shared-memory instruction.  byte_addresses: 32 lane addresses."""
    words = set()
    for a in byte_addresses:
        for k in range(access_bytes // 4):
            words.add(int(a) // 4 + k)
    per_bank = np.bincount([w % 32 for w in words], minlength=32)
    return int(per_bank.max()), -(-len(words) * 4 // 128)


def column_tile_wavefronts(N=512, W=4):
    """LSU wavefronts per tile of the TMA column kernel by instruction class: {class: (wavefronts, ideal)}.
    Thread = (tid, xw), lane = xw + W * tid; element e of a tile buffer sits at byte 16 * e (exchange 0 dense, exchange
    1.. swizzled); twiddles are float2 [(stage*7 + q-1) * T + tid]."""
    T, S = N // 8, len(stage_radices(N))
    threads = W * T
    out = {}

    def add(kind, addr_fn, nbytes, count=1):
        for w0 in range(0, threads, 32):
            lanes = range(w0, w0 + 32)
            wf, ideal = warp_wavefronts([addr_fn(l // W, l % W) for l in lanes], nbytes)
            a = out.setdefault(kind, [0, 0])
            a[0] += wf * count
            a[1] += ideal * count

    for plane in range(2):
        for j in range(8):  # initial loads of the TMA-staged tile and the final staging stores: element (tid + T*j) * W + xw
            add("LDS.128", lambda tid, xw, j=j: 16 * ((tid + T * j) * W + xw), 16)
            add("STS.128", lambda tid, xw, j=j: 16 * ((tid + T * j) * W + xw), 16)
        for s in range(S - 1):
            Cs = 8 ** s
            Ms = N // Cs
            Mn = Ms // 8
            for q in range(8):  # exchange stores: logical F = tid + T*q (ocean_fft.cuh stage<s>)
                add("STS.128", lambda tid, xw, q=q, s=s: 16 * swizzle((tid + T * q) * W + xw, s, True), 16)
            rn = stage_radices(N)[s + 1]
            for j in range(8):
                if rn == 8:
                    Q = Mn // 8
                    add("LDS.128", lambda tid, xw, j=j, s=s, Q=Q: 16 * swizzle((((tid // Q) * 8 + j) * Q + tid % Q) * W + xw, s, True), 16)
                else:
                    i, jj = divmod(j, rn)
                    add("LDS.128", lambda tid, xw, i=i, jj=jj, s=s, rn=rn: 16 * swizzle(((tid + T * i) * rn + jj) * W + xw, s, True), 16)
            for q in range(1, 8):  # twiddles of the stage that precedes this exchange
                add("LDS.64", lambda tid, xw, q=q, s=s: 8 * ((s * 7 + q - 1) * T + tid), 8)
    for j in range(8):  # plane B's epilogue: incoming foam = .x of the staged RGBA turbulence texel, outgoing texel
        add("LDS.32", lambda tid, xw, j=j: 16 * ((tid + T * j) * W + xw), 4)
        add("STS.128", lambda tid, xw, j=j: 16 * ((tid + T * j) * W + xw), 16)
    return {k: tuple(v) for k, v in out.items()}
