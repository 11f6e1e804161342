#!/usr/bin/env python
"""Emit straight-line in-register radix-2 DIT FFT code (N = 16, 32) for the warp-level FFT.

Input is expected in bit-reversed register order, output is in natural order.  Every butterfly
is written with explicit FMAs (Linzer-Feig form: X = a + w*b with two nested FMAs per component,
Y = 2a - X with one), so a general butterfly costs 6 FP32 instructions, a +-(1+-i)/sqrt(2)
butterfly 6, and a trivial (1, -i) butterfly 4.

    python tools/gen_fft.py > audio_feature_extraction_b200/csrc/fft_gen.cuh
"""
import math


def bitrev(n, bits):
    return int(format(n, "0%db" % bits)[::-1], 2)


def emit_fft(N, real_t="float", copies=1):
    bits = N.bit_length() - 1
    out = []
    suffix = "f" if real_t == "float" else ""
    fma = "fmaf" if real_t == "float" else "fma"
    out.append("// %d-point DIT, bit-reversed in -> natural out; forward transform exp(-2*pi*i*nk/N)" % N)
    if copies > 1:
        out.append("// %d independent transforms on consecutive blocks of %d registers" % (copies, N))
        out.append("__device__ __forceinline__ void fft%dx%d_dit(%s (&r)[%d], %s (&i)[%d]) {" % (N, copies, real_t, N * copies, real_t, N * copies))
    else:
        out.append("__device__ __forceinline__ void fft%d_dit(%s (&r)[%d], %s (&i)[%d]) {" % (N, real_t, N, real_t, N))
    out.append("  %s xr, xi, s, d;" % real_t)
    n_instr = 0
    L = 2
    while L <= N:
        h = L // 2
        out.append("  // stage len=%d" % L)
        for base in range(0, N * copies, L):
            for j in range(h):
                p, q = base + j, base + j + h
                # twiddle index on the unit circle in 1/N' turns where N' = L
                num, den = j, L
                g = math.gcd(num, den) if num else den
                num //= g
                den //= g
                if num == 0:
                    out.append("  xr = r[%d]; xi = i[%d]; r[%d] = r[%d] - xr; i[%d] = i[%d] - xi; r[%d] += xr; i[%d] += xi;"
                               % (q, q, q, p, q, p, p, p))
                    n_instr += 4
                elif (num, den) == (1, 4):  # w = -i : w*b = (b.i, -b.r)
                    out.append("  xr = i[%d]; xi = r[%d]; r[%d] = r[%d] - xr; i[%d] = i[%d] + xi; r[%d] += xr; i[%d] -= xi;"
                               % (q, q, q, p, q, p, p, p))
                    n_instr += 4
                elif (num, den) == (1, 8):  # w = c(1 - i): w*b = c((br+bi) + i(bi-br))
                    c = repr(math.sqrt(0.5)) + suffix
                    out.append("  s = r[%d] + i[%d]; d = i[%d] - r[%d];" % (q, q, q, q))
                    out.append("  r[%d] = %s(-%s, s, r[%d]); i[%d] = %s(-%s, d, i[%d]); r[%d] = %s(%s, s, r[%d]); i[%d] = %s(%s, d, i[%d]);"
                               % (q, fma, c, p, q, fma, c, p, p, fma, c, p, p, fma, c, p))
                    n_instr += 6
                elif (num, den) == (3, 8):  # w = -c(1 + i): w*b = -c((br-bi) + i(br+bi))
                    c = repr(math.sqrt(0.5)) + suffix
                    out.append("  s = r[%d] - i[%d]; d = r[%d] + i[%d];" % (q, q, q, q))
                    out.append("  r[%d] = %s(%s, s, r[%d]); i[%d] = %s(%s, d, i[%d]); r[%d] = %s(-%s, s, r[%d]); i[%d] = %s(-%s, d, i[%d]);"
                               % (q, fma, c, p, q, fma, c, p, p, fma, c, p, p, fma, c, p))
                    n_instr += 6
                else:
                    ang = 2.0 * math.pi * j / L
                    wr = repr(math.cos(ang)) + suffix
                    wi = repr(math.sin(ang)) + suffix  # w = wr - i*wi
                    # X = a + w b:  Xr = ar + wr*br + wi*bi ; Xi = ai + wr*bi - wi*br
                    out.append("  xr = %s(%s, r[%d], %s(%s, i[%d], r[%d])); xi = %s(%s, i[%d], %s(-%s, r[%d], i[%d]));"
                               % (fma, wr, q, fma, wi, q, p, fma, wr, q, fma, wi, q, p))
                    out.append("  r[%d] = %s(2.0%s, r[%d], -xr); i[%d] = %s(2.0%s, i[%d], -xi); r[%d] = xr; i[%d] = xi;"
                               % (q, fma, suffix, p, q, fma, suffix, p, p, p))
                    n_instr += 6
        L *= 2
    out.append("}  // ~%d FP32 instructions" % n_instr)
    return "\n".join(out)


def emit_fft_packed(N, copies=1):
    """Same transform on (re, im) pairs packed in 64-bit registers: FADD2 / FFMA2 (sm_100a) do both components in one
    issue slot; swaps and sign flips of the halves are operand modifiers, so a trivial or -i butterfly costs 2
    instructions, a (1 -+ i)/sqrt(2) or general one 3."""
    out = []
    out.append("// %d-point DIT on packed (re, im) pairs, bit-reversed in -> natural out; forward transform" % N)
    if copies > 1:
        out.append("// %d independent transforms on consecutive blocks of %d registers" % (copies, N))
        out.append("__device__ __forceinline__ void fft%dx%d_dit2(c64 (&z)[%d]) {" % (N, copies, N * copies))
    else:
        out.append("__device__ __forceinline__ void fft%d_dit2(c64 (&z)[%d]) {" % (N, N))
    out.append("  c64 t, x;")
    n_instr = 0
    L = 2
    while L <= N:
        h = L // 2
        out.append("  // stage len=%d" % L)
        for base in range(0, N * copies, L):
            for j in range(h):
                p, q = base + j, base + j + h
                num, den = j, L
                g = math.gcd(num, den) if num else den
                num //= g
                den //= g
                if num == 0:
                    out.append("  t = z[%d]; z[%d] = sub2(z[%d], t); z[%d] = add2(z[%d], t);" % (q, q, p, p, p))
                    n_instr += 2
                elif (num, den) == (1, 4):  # w b = (b.i, -b.r)
                    out.append("  t = mul_mi(z[%d]); z[%d] = sub2(z[%d], t); z[%d] = add2(z[%d], t);" % (q, q, p, p, p))
                    n_instr += 2
                elif (num, den) == (1, 8):  # w b = c (b + (-i) b)
                    c = repr(math.sqrt(0.5)) + "f"
                    out.append("  t = add2(z[%d], mul_mi(z[%d])); z[%d] = fma2(t, bc2(-%s), z[%d]); z[%d] = fma2(t, bc2(%s), z[%d]);"
                               % (q, q, q, c, p, p, c, p))
                    n_instr += 3
                elif (num, den) == (3, 8):  # w b = -c (b + i b)
                    c = repr(math.sqrt(0.5)) + "f"
                    out.append("  t = add2(z[%d], mul_pi(z[%d])); z[%d] = fma2(t, bc2(%s), z[%d]); z[%d] = fma2(t, bc2(-%s), z[%d]);"
                               % (q, q, q, c, p, p, c, p))
                    n_instr += 3
                else:
                    ang = 2.0 * math.pi * j / L
                    wr = repr(math.cos(ang)) + "f"
                    wi = repr(math.sin(ang)) + "f"  # w = wr - i wi: w b = wr b + wi (b.i, -b.r)
                    out.append("  x = fma2(swap2(z[%d]), pk2(%s, -%s), fma2(z[%d], bc2(%s), z[%d])); z[%d] = fma2(z[%d], bc2(2.0f), neg2(x)); z[%d] = x;"
                               % (q, wi, wi, q, wr, p, q, p, p))
                    n_instr += 3
        L *= 2
    out.append("}  // ~%d packed FP32 instructions" % n_instr)
    return "\n".join(out)


PACKED_PRELUDE = """// ---- packed FP32 pairs (sm_100a FADD2 / FMUL2 / FFMA2): one issue slot for both components of a complex value;
// ptxas folds the swaps / sign flips built with pk2() into operand modifiers and broadcast constants into immediates
typedef unsigned long long c64;   // (re = low half, im = high half)
__device__ __forceinline__ c64 pk2(float lo, float hi) { c64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo2(c64 v) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(v)); (void)h; return l; }
__device__ __forceinline__ float hi2(c64 v) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(v)); (void)l; return h; }
__device__ __forceinline__ c64 bc2(float c) { return pk2(c, c); }
__device__ __forceinline__ c64 swap2(c64 v) { return pk2(hi2(v), lo2(v)); }
__device__ __forceinline__ c64 neg2(c64 v) { return pk2(-lo2(v), -hi2(v)); }
__device__ __forceinline__ c64 mul_mi(c64 v) { return pk2(hi2(v), -lo2(v)); }   // -i v
__device__ __forceinline__ c64 mul_pi(c64 v) { return pk2(-hi2(v), lo2(v)); }   // +i v
__device__ __forceinline__ c64 add2(c64 a, c64 b) { c64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c64 sub2(c64 a, c64 b) { c64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c64 mul2(c64 a, c64 b) { c64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c64 fma2(c64 a, c64 b, c64 c) { c64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
// complex product v * (w.x + i w.y)
__device__ __forceinline__ c64 cmul2(c64 v, float wx, float wy) { return fma2(swap2(v), pk2(-wy, wy), mul2(v, bc2(wx))); }
"""


def main():
    print("// GENERATED by tools/gen_fft.py -- do not edit by hand.")
    print("#pragma once")
    print("namespace afx {")
    for N in (16, 32):
        bits = N.bit_length() - 1
        print("__device__ __forceinline__ constexpr int bitrev%d(int n) { return %s; }"
              % (N, " | ".join("(((n >> %d) & 1) << %d)" % (b, bits - 1 - b) for b in range(bits))))
        print(emit_fft(N))
        print()
    print(emit_fft(16, copies=2))
    print()
    print(PACKED_PRELUDE)
    print(emit_fft_packed(32))
    print()
    print(emit_fft_packed(16, copies=2))
    print()
    print("}  // namespace afx")


if __name__ == "__main__":
    main()
