// Empirical check: for divisors d whose reciprocal y = RN(1/d) is precomputed, is
//   q = a*y;  r = fma(-d, q, a);  q' = fma(r, y, q)
// bit-identical to the IEEE quotient a/d (__ddiv_rn)?  Counts mismatches over many random
// and adversarial numerators.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t splitmix(uint64_t &s)
{
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

template <int STEPS>
__device__ __forceinline__ double fast_div(double a, double d, double y)
{
    double q = a * y;
#pragma unroll
    for (int s = 0; s < STEPS; ++s) {
        double r = fma(-d, q, a);
        q = fma(r, y, q);
    }
    return q;
}

template <int STEPS>
__global__ void check(const double *ds, const double *ys, int nd, int iters, uint64_t seed,
                      unsigned long long *mismatch, double *example)
{
    uint64_t s = seed + (uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0x1234567ull;
    unsigned long long bad = 0;
    for (int it = 0; it < iters; ++it) {
        uint64_t u = splitmix(s);
        const int di = (int)(splitmix(s) % (uint64_t)nd);
        const double d = ds[di], y = ys[di];
        // numerator: random mantissa; exponent in [-60, 60]; every 4th sample is built to be
        // close to a quotient midpoint: a = RN(d * (m + 1/2 ulp))
        uint64_t mant = u & 0xfffffffffffffull;
        int ex = (int)((u >> 52) % 121) - 60;
        double a = __longlong_as_double((long long)(((uint64_t)(1023 + ex) << 52) | mant));
        if (u >> 63) a = -a;
        if ((it & 3) == 3) {
            double m = __longlong_as_double((long long)((1023ull << 52) | mant));
            double half = ldexp(1.0, -53);
            a = d * (m + ((u >> 62) & 1 ? half : -half) * (1.0 + ldexp((double)(u & 255), -60)));
        }
        const double ref = __ddiv_rn(a, d);
        const double got = fast_div<STEPS>(a, d, y);
        if (__double_as_longlong(ref) != __double_as_longlong(got)) {
            if (bad == 0 && atomicAdd(mismatch + 1, 1ull) < 4) { example[0] = a; example[1] = d; }
            ++bad;
        }
    }
    if (bad) atomicAdd(mismatch, bad);
}

// Second rule (csrc/mpx_dds.cu): for d = s 2^j (1 - 2^-53) -- significand all ones, where the
// sequence above is NOT exact -- the quotient is the next double after A = a / (s 2^j), away from
// zero (for a != 0).  Computed exactly as the kernel does: the three instructions with the pair
// (s 2^j, s 2^-j), then one integer increment of the bit pattern.
__global__ void check_all_ones(int iters, uint64_t seed, unsigned long long *mismatch, double *example)
{
    uint64_t s = seed + (uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0x7654321ull;
    unsigned long long bad = 0;
    for (int it = 0; it < iters; ++it) {
        uint64_t u = splitmix(s);
        const int j = (int)(splitmix(s) % 401) - 200;                 // 2^j, |j| <= 200
        const double sgn = (u >> 61) & 1 ? -1.0 : 1.0;
        const double pw = sgn * ldexp(1.0, j);
        const double d = sgn * (ldexp(1.0, j) - ldexp(1.0, j - 53));  // all-ones significand
        uint64_t mant = u & 0xfffffffffffffull;
        if ((it & 7) == 7) mant = 0;                                  // powers of two: the tie case
        if ((it & 7) == 6) mant = 0xfffffffffffffull;                 // top of the binade
        if ((it & 7) == 5) mant &= 0xffull;                           // just above a power of two
        int ex = (int)((u >> 52) % 1401) - 700;                       // 2^-700 .. 2^700
        double a = __longlong_as_double((long long)(((uint64_t)(1023 + ex) << 52) | mant));
        if (u >> 63) a = -a;
        const double ref = __ddiv_rn(a, d);
        double q = fast_div<1>(a, pw, 1.0 / pw);
        q = __longlong_as_double(__double_as_longlong(q) + 1);
        if (__double_as_longlong(ref) != __double_as_longlong(q)) {
            if (bad == 0 && atomicAdd(mismatch + 1, 1ull) < 4) { example[0] = a; example[1] = d; }
            ++bad;
        }
    }
    if (bad) atomicAdd(mismatch, bad);
}

int main()
{
    // divisors: all pairwise differences of Chebyshev-Lobatto points (any order gives the same
    // set as Leja ordering) for n = 1..64, plus pseudo-random ones
    std::vector<double> ds;
    for (int n = 1; n <= 64; ++n) {
        std::vector<double> p(n + 1);
        for (int i = 0; i <= n; ++i) p[i] = std::cos(M_PI * i / n);
        for (int i = 0; i <= n; ++i)
            for (int j = 0; j <= n; ++j)
                if (i != j) ds.push_back(p[i] - p[j]);
    }
    uint64_t s = 42;
    for (int i = 0; i < 200000; ++i) {
        s = s * 6364136223846793005ull + 1442695040888963407ull;
        double m = 1.0 + (double)(s >> 12) / 4503599627370496.0;
        int ex = (int)((s >> 3) % 41) - 20;
        ds.push_back(std::ldexp((s & 1) ? m : -m, ex));
    }
    std::vector<double> ys(ds.size());
    for (size_t i = 0; i < ds.size(); ++i) ys[i] = 1.0 / ds[i];
    double *dd, *dy, *dex;
    unsigned long long *dm;
    cudaMalloc(&dd, ds.size() * 8); cudaMalloc(&dy, ds.size() * 8); cudaMalloc(&dm, 16); cudaMalloc(&dex, 16);
    cudaMemcpy(dd, ds.data(), ds.size() * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(dy, ys.data(), ds.size() * 8, cudaMemcpyHostToDevice);
    const int blocks = 148 * 16, threads = 256, iters = 1 << 14;
    for (int steps = 1; steps <= 2; ++steps) {
        cudaMemset(dm, 0, 16);
        if (steps == 1) check<1><<<blocks, threads>>>(dd, dy, (int)ds.size(), iters, 777, dm, dex);
        else check<2><<<blocks, threads>>>(dd, dy, (int)ds.size(), iters, 777, dm, dex);
        cudaDeviceSynchronize();
        unsigned long long m[2]; double ex[2];
        cudaMemcpy(m, dm, 16, cudaMemcpyDeviceToHost);
        cudaMemcpy(ex, dex, 16, cudaMemcpyDeviceToHost);
        printf("{\"correction_steps\": %d, \"divisors\": %zu, \"samples\": %.3e, \"mismatches\": %llu, \"example_a\": %.17g, \"example_d\": %.17g}\n",
               steps, ds.size(), (double)blocks * threads * iters, m[0], m[0] ? ex[0] : 0.0, m[0] ? ex[1] : 0.0);
    }
    {
        cudaMemset(dm, 0, 16);
        check_all_ones<<<blocks, threads>>>(iters, 4242, dm, dex);
        cudaDeviceSynchronize();
        unsigned long long m[2]; double ex[2];
        cudaMemcpy(m, dm, 16, cudaMemcpyDeviceToHost);
        cudaMemcpy(ex, dex, 16, cudaMemcpyDeviceToHost);
        printf("{\"rule\": \"all-ones significand: next double away from zero\", \"samples\": %.3e, \"mismatches\": %llu, \"example_a\": %.17g, \"example_d\": %.17g}\n",
               (double)blocks * threads * iters, m[0], m[0] ? ex[0] : 0.0, m[0] ? ex[1] : 0.0);
    }
    return 0;
}
