"""Generates ditsep_b200/csrc/dft_codelets.cuh: straight-line 15- and 17-point complex DFT codelets with literal
twiddle constants, the two factors of the 255-point transform behind the 510-point real STFT (csrc/spectral.cu).

    python scripts/gen_dft_codelets.py            # rewrites the header
    python scripts/gen_dft_codelets.py --check    # g++ host harness: codelets vs a double-precision naive DFT

15 = 3 x 5 with coprime factors: Good-Thomas index maps (n = 5 n1 + 3 n2, k = 10 k1 + 6 k2 mod 15) need no twiddles
between the 3- and 5-point butterflies.  17 is prime: inputs n and 17 - n are paired, a_n = x_n + x_{17-n},
b_n = x_n - x_{17-n}, so X[k], X[17-k] = x_0 + sum_n a_n cos(2 pi n k / 17) -/+ i sum_n b_n sin(2 pi n k / 17): two 8 x 8
real-coefficient products on complex data instead of a 17 x 17 complex one.  Forward sign e^{-i...}; the inverse is the
same codelet on inputs with real and imaginary parts swapped (done by the caller)."""
import math
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "ditsep_b200", "csrc", "dft_codelets.cuh")


def f(v):
    return f"{v:.9e}f"


def gen_dft3(a, b, c, tag):
    """In-place forward 3-point DFT on the float2 lvalues a, b, c."""
    s = math.sin(2 * math.pi / 3)
    L = [f"  {{  // dft3 {tag}",
         f"    const float tx = {b}.x + {c}.x, ty = {b}.y + {c}.y;",
         f"    const float ux = {a}.x - 0.5f * tx, uy = {a}.y - 0.5f * ty;",
         f"    const float vx = {f(s)} * ({b}.y - {c}.y), vy = {f(-s)} * ({b}.x - {c}.x);",  # -i s (b - c)
         f"    {a}.x += tx; {a}.y += ty;",
         f"    {b}.x = ux + vx; {b}.y = uy + vy;",
         f"    {c}.x = ux - vx; {c}.y = uy - vy;",
         "  }"]
    return L


def gen_dft5(v, tag):
    """In-place forward 5-point DFT on v[0..4] (names of float2 lvalues)."""
    c1, c2 = math.cos(2 * math.pi / 5), math.cos(4 * math.pi / 5)
    s1, s2 = math.sin(2 * math.pi / 5), math.sin(4 * math.pi / 5)
    a, b, c, d, e = v
    L = [f"  {{  // dft5 {tag}",
         f"    const float p1x = {b}.x + {e}.x, p1y = {b}.y + {e}.y, m1x = {b}.x - {e}.x, m1y = {b}.y - {e}.y;",
         f"    const float p2x = {c}.x + {d}.x, p2y = {c}.y + {d}.y, m2x = {c}.x - {d}.x, m2y = {c}.y - {d}.y;",
         f"    const float r1x = {a}.x + {f(c1)} * p1x + {f(c2)} * p2x, r1y = {a}.y + {f(c1)} * p1y + {f(c2)} * p2y;",
         f"    const float r2x = {a}.x + {f(c2)} * p1x + {f(c1)} * p2x, r2y = {a}.y + {f(c2)} * p1y + {f(c1)} * p2y;",
         # q_k = sum_n m_n sin(2 pi n k / 5): q1 = s1 m1 + s2 m2, q2 = s2 m1 - s1 m2;  X[k] = r_k - i q_k, X[5-k] = r_k + i q_k
         f"    const float q1x = {f(s1)} * m1x + {f(s2)} * m2x, q1y = {f(s1)} * m1y + {f(s2)} * m2y;",
         f"    const float q2x = {f(s2)} * m1x - {f(s1)} * m2x, q2y = {f(s2)} * m1y - {f(s1)} * m2y;",
         f"    {a}.x += p1x + p2x; {a}.y += p1y + p2y;",
         f"    {b}.x = r1x + q1y; {b}.y = r1y - q1x;",
         f"    {e}.x = r1x - q1y; {e}.y = r1y + q1x;",
         f"    {c}.x = r2x + q2y; {c}.y = r2y - q2x;",
         f"    {d}.x = r2x - q2y; {d}.y = r2y + q2x;",
         "  }"]
    return L


def gen_dft15():
    L = ["// out[k] = sum_n in[n] e^{-2 pi i n k / 15}",
         "DFT_HD void dft15(const float2 (&in)[15], float2 (&out)[15]) {",
         "  float2 t[3][5];  // t[n1][n2] = in[(5 n1 + 3 n2) mod 15]"]
    for n1 in range(3):
        for n2 in range(5):
            L.append(f"  t[{n1}][{n2}] = in[{(5 * n1 + 3 * n2) % 15}];")
    for n2 in range(5):
        L += gen_dft3(f"t[0][{n2}]", f"t[1][{n2}]", f"t[2][{n2}]", f"column {n2}")
    for k1 in range(3):
        L += gen_dft5([f"t[{k1}][{n2}]" for n2 in range(5)], f"row {k1}")
    for k1 in range(3):
        for k2 in range(5):
            L.append(f"  out[{(10 * k1 + 6 * k2) % 15}] = t[{k1}][{k2}];")
    L.append("}")
    return L


def gen_dft17():
    L = ["// out[k] = sum_n in[n] e^{-2 pi i n k / 17}",
         "DFT_HD void dft17(const float2 (&in)[17], float2 (&out)[17]) {",
         "  float2 a[9], b[9];"]
    for n in range(1, 9):
        L.append(f"  a[{n}] = make_float2(in[{n}].x + in[{17 - n}].x, in[{n}].y + in[{17 - n}].y);")
        L.append(f"  b[{n}] = make_float2(in[{n}].x - in[{17 - n}].x, in[{n}].y - in[{17 - n}].y);")
    L.append("  out[0] = make_float2(in[0].x + " + " + ".join(f"a[{n}].x" for n in range(1, 9)) + ",")
    L.append("                       in[0].y + " + " + ".join(f"a[{n}].y" for n in range(1, 9)) + ");")
    for k in range(1, 9):
        cx = " + ".join(f"{f(math.cos(2 * math.pi * ((n * k) % 17) / 17))} * a[{n}].x" for n in range(1, 9))
        cy = " + ".join(f"{f(math.cos(2 * math.pi * ((n * k) % 17) / 17))} * a[{n}].y" for n in range(1, 9))
        sx = " + ".join(f"{f(math.sin(2 * math.pi * ((n * k) % 17) / 17))} * b[{n}].x" for n in range(1, 9))
        sy = " + ".join(f"{f(math.sin(2 * math.pi * ((n * k) % 17) / 17))} * b[{n}].y" for n in range(1, 9))
        L += ["  {",
              f"    const float cx = in[0].x + {cx};",
              f"    const float cy = in[0].y + {cy};",
              f"    const float sx = {sx};",
              f"    const float sy = {sy};",
              f"    out[{k}] = make_float2(cx + sy, cy - sx);",       # c - i s
              f"    out[{17 - k}] = make_float2(cx - sy, cy + sx);",  # c + i s
              "  }"]
    L.append("}")
    return L


HEADER = """// GENERATED by scripts/gen_dft_codelets.py -- do not edit; regenerate instead.
// Straight-line 15- and 17-point forward complex DFTs with literal twiddles (the factors of the 255-point transform
// of csrc/spectral.cu).  Inverse transforms: swap .x / .y of the inputs and of the outputs.
#pragma once
#if defined(__CUDACC__)
#define DFT_HD __host__ __device__ __forceinline__
#else
#define DFT_HD inline
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
#endif

namespace ditsep {
"""


def generate():
    lines = [HEADER.rstrip("\n"), ""] + gen_dft15() + [""] + gen_dft17() + ["", "}  // namespace ditsep", ""]
    return "\n".join(lines)


CHECK = r"""
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "dft_codelets.cuh"
template <int N, typename F> double check(F fn) {
  float2 in[N], out[N];
  double worst = 0;
  for (int trial = 0; trial < 50; ++trial) {
    for (int n = 0; n < N; ++n) in[n] = make_float2((float)rand() / RAND_MAX - 0.5f, (float)rand() / RAND_MAX - 0.5f);
    fn(in, out);
    for (int k = 0; k < N; ++k) {
      double re = 0, im = 0;
      for (int n = 0; n < N; ++n) {
        double a = -2.0 * M_PI * ((n * k) % N) / N;
        re += in[n].x * cos(a) - in[n].y * sin(a);
        im += in[n].x * sin(a) + in[n].y * cos(a);
      }
      worst = fmax(worst, fmax(fabs(re - out[k].x), fabs(im - out[k].y)));
    }
  }
  return worst;
}
int main() {
  double e15 = check<15>([](const float2 (&a)[15], float2 (&b)[15]) { ditsep::dft15(a, b); });
  double e17 = check<17>([](const float2 (&a)[17], float2 (&b)[17]) { ditsep::dft17(a, b); });
  printf("dft15 max abs err %.3e, dft17 max abs err %.3e\n", e15, e17);
  return (e15 < 5e-6 && e17 < 5e-6) ? 0 : 1;
}
"""


def main():
    text = generate()
    if "--check" in sys.argv:
        with tempfile.TemporaryDirectory() as d:
            open(os.path.join(d, "dft_codelets.cuh"), "w").write(text)
            open(os.path.join(d, "check.cpp"), "w").write(CHECK)
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", d, os.path.join(d, "check.cpp"), "-o", os.path.join(d, "check")])
            sys.exit(subprocess.call([os.path.join(d, "check")]))
    open(OUT, "w").write(text)
    print("written", OUT, f"({len(text.splitlines())} lines)")


if __name__ == "__main__":
    main()
