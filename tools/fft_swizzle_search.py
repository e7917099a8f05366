"""Shared-memory bank model of k_dct1_fft's line buffer (csrc/gdpt_poisson_fft.cuh) and the search that produced fft_sw.

An element is a double2 (16 bytes), so a warp's access is served a quarter-warp at a time over 8 sixteen-byte banks; a quarter-warp
costs as many wavefronts as its most loaded bank.  passes(P, logP) enumerates the index patterns of every access of fft_forward /
fft_inverse (radix-4 passes, the radix-2 pass of an odd log2 P) and of the contiguous sweeps; cost() = wavefronts / ideal.
Run as a script: searches all maps  low3(e) ^= M * bits 3..6 of e  (M: four 3-bit masks) and prints the conflict-free ones.
tests/test_host_cpu.py asserts that the map in the header (masks 3 and 6 on bits 3 and 4: x ^ x<<1 with x = bits 3..4) is one."""
import itertools


def fft_sw(e):
    x = (e >> 3) & 3
    return e ^ (x ^ (x << 1))


def wavefronts(addrs):
    tot = 0
    for q in range(4):
        banks = {}
        for a in addrs[8 * q:8 * q + 8]:
            banks.setdefault(a % 8, set()).add(a)
        tot += max(len(v) for v in banks.values())
    return tot


def passes(P, logP):
    pats = []

    def radix4(n, ln, tag):
        q4, lq = n >> 2, ln - 2
        for k in range(4):
            pats.append(("%s q4=%d k=%d" % (tag, q4, k), lambda idx, q4=q4, lq=lq, ln=ln, k=k: ((idx >> lq) << ln) + (idx & (q4 - 1)) + k * q4))
    n, ln = P, logP
    for _ in range(logP // 2):
        radix4(n, ln, "fwd")
        n >>= 2; ln -= 2
    if logP & 1:
        for c in range(2):
            pats.append(("radix-2 c=%d" % c, lambda idx, c=c: 2 * idx + c))
        n, ln = 8, 3
        for _ in range(logP // 2):
            radix4(n, ln, "inv")
            n <<= 2; ln += 2
    pats.append(("contiguous", lambda idx: idx))
    return pats


def cost(swz, P, logP):
    tot = ideal = 0
    worst = {}
    for name, f in passes(P, logP):
        for warp in range(4):
            w = wavefronts([swz(f(warp * 32 + lane)) for lane in range(32)])
            tot += w; ideal += 4
            worst[name] = max(worst.get(name, 0), w)
    return tot / ideal, worst


SIZES = ((2048, 11), (4096, 12), (8192, 13))

if __name__ == "__main__":
    for P, l in SIZES:
        print("P = %d: natural layout %.2f x the ideal wavefronts, fft_sw %.2f x" % (P, cost(lambda e: e, P, l)[0], cost(fft_sw, P, l)[0]))
    good = []
    for m in itertools.product(range(8), repeat=4):
        swz = lambda e, m=m: e ^ ((m[0] if e >> 3 & 1 else 0) ^ (m[1] if e >> 4 & 1 else 0) ^ (m[2] if e >> 5 & 1 else 0) ^ (m[3] if e >> 6 & 1 else 0))
        if all(cost(swz, P, l)[0] == 1.0 for P, l in SIZES):
            good.append(m)
    print("%d conflict-free maps; masks on bits 3, 4 (bits 5, 6 free):" % len(good), sorted(set(m[:2] for m in good)))
