// pt_post.cuh — what the reference's main() does after the sample loop (main.cpp:162-201), on the device:
// dispersion statistics + gamma resolve, the optional Gauss / median filter (main.cpp:11-33, 49-80) and the 8-bit
// conversion of bitmap_image::set_pixel, so that 3 bytes per pixel leave the GPU instead of 28.
//
// Exactness: everything here reproduces the reference bit for bit.  pow() is glibc's powf restated (ref_powf);
// the Gaussian weights depend only on the integer offset and are tabulated on the HOST with the reference's own
// expression (expf of the host libm); sums run in the reference's order; the average dispersion is a sequential fp32
// sum over the pixels in the reference's loop order (k_dispersion_sum: one thread adds, the block streams).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "pt_libm.cuh"

namespace ptd {

// glibc powf (sysdeps/ieee754/flt-32/e_powf.c, the FMA build that x86-64 selects) for x >= 0 and finite y > 0:
// log2 by a 16-entry table + degree-5 polynomial in double, exp2 by a 32-entry table + degree-3 polynomial, one
// rounding to float.  Bit-equal to the host's powf (tests: oracle restatement on 10^8 inputs, device on 2^24).
__device__ __forceinline__ float ref_powf(float x, float y) {
    uint32_t ix = __float_as_uint(x);
    if (x != x || y != y) return x + y;
    if (ix & 0x80000000u) return ix == 0x80000000u ? 0.0f : __int_as_float(0x7fc00000);
    if (ix == 0u) return 0.0f;
    if (ix == 0x7f800000u) return x;
    if (ix < 0x00800000u) {  // subnormal: normalise
        ix = __float_as_uint(x * 8388608.0f);
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (tmp >> 19) & 15;
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz = ix - top;
    const int k = (int)top >> 23;
    const double invc = kPowLog2Tab[2 * i], logc = kPowLog2Tab[2 * i + 1];
    const double z = (double)__uint_as_float(iz);
    const double r = __fma_rn(z, invc, -1.0);
    const double y0 = __dadd_rn(logc, (double)k);
    const double r2 = __dmul_rn(r, r);
    double yy = __fma_rn(kPowLog2Poly[0], r, kPowLog2Poly[1]);
    const double p = __fma_rn(kPowLog2Poly[2], r, kPowLog2Poly[3]);
    const double r4 = __dmul_rn(r2, r2);
    double q = __fma_rn(kPowLog2Poly[4], r, y0);
    q = __fma_rn(p, r2, q);
    yy = __fma_rn(yy, r4, q);
    const double ylogx = __dmul_rn((double)y, yy);
    if (((__double_as_longlong(ylogx) >> 47) & 0xffff) >= (__double_as_longlong(126.0) >> 47)) {
        if (ylogx > 0x1.fffffffd1d571p+6) return __int_as_float(0x7f800000);
        if (ylogx <= -150.0) return 0.0f;
    }
    const double shift = 0x1.8p+47;
    double kd = __dadd_rn(ylogx, shift);
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd = __dadd_rn(kd, -shift);
    const double rr = __dadd_rn(ylogx, -kd);
    unsigned long long t = kExp2fTab[ki & 31];
    t += ki << 47;
    const double sc = __longlong_as_double((long long)t);
    const double zz = __fma_rn(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3);
    const double rr2 = __dmul_rn(rr, rr);
    double e = __fma_rn(0x1.62e42ff0c52d6p-1, rr, 1.0);
    e = __fma_rn(zz, rr2, e);
    e = __dmul_rn(e, sc);
    return (float)e;
}

__global__ void k_powf(const float *__restrict__ x, long long n, float y, float *__restrict__ out) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < n) out[i] = ref_powf(x[i], y);
}

// main.cpp:162-185 per pixel.  color = pow(sum / n, gamma) * 255 where the pixel has samples, else the untouched sum;
// disp = D.r + D.g + D.b with D = sum2 / n - (sum / n)^2, or the 1 an unsampled pixel adds to the average.
// minmax[0] / [1]: ordered-int max / min of disp over the sampled pixels (initialised by the caller to 0 / +inf).
__global__ void __launch_bounds__(256)
k_tonemap(const float *__restrict__ sum3, const float *__restrict__ sum2_3, const int *__restrict__ count,
          const float *__restrict__ weight, long long npx, float gamma, float *__restrict__ color3,
          float *__restrict__ disp, int *minmax) {
    const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    float dmax = 0.0f, dmin = __int_as_float(0x7f800000);
    if (p < npx) {
        const float sx = sum3[3 * p], sy = sum3[3 * p + 1], sz = sum3[3 * p + 2];
        if (!count[p]) {
            color3[3 * p] = sx; color3[3 * p + 1] = sy; color3[3 * p + 2] = sz;
            disp[p] = 1.0f;
        } else {
            const float n = weight ? weight[p] : (float)count[p];
            const float mx = sx / n, my = sy / n, mz = sz / n;
            const float Dx = sum2_3[3 * p] / n - mx * mx, Dy = sum2_3[3 * p + 1] / n - my * my,
                        Dz = sum2_3[3 * p + 2] / n - mz * mz;
            const float dd = Dx + Dy + Dz;
            disp[p] = dd;
            if (dd > dmax) dmax = dd;  // the reference's comparisons: a NaN never wins
            if (dd < dmin) dmin = dd;
            color3[3 * p] = ref_powf(mx, gamma) * 255.0f;
            color3[3 * p + 1] = ref_powf(my, gamma) * 255.0f;
            color3[3 * p + 2] = ref_powf(mz, gamma) * 255.0f;
        }
    }
    for (int o = 16; o; o >>= 1) {
        dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
        dmin = fminf(dmin, __shfl_xor_sync(0xffffffffu, dmin, o));
    }
    if ((threadIdx.x & 31) == 0) {  // dmax >= 0: its bits order as ints; dmin may be negative: ordered-int encoding
        atomicMax(&minmax[0], __float_as_int(dmax));
        const int b = __float_as_int(dmin);
        atomicMin(&minmax[1], b >= 0 ? b : b ^ 0x7fffffff);
    }
}

// average_dispersion (main.cpp:165-179): a sequential fp32 sum over the pixels in the loop order y outer, x inner.
// fp32 addition is not associative, so ONE thread adds in that order; the other threads of the block stream the
// next chunk into shared memory meanwhile (double buffer).  ~4 cycles per pixel: 4 ms at 1080p.
constexpr int kSeqChunk = 4096;
__global__ void __launch_bounds__(1024) k_dispersion_sum(const float *__restrict__ disp, long long npx, float *out) {
    __shared__ float buf[2][kSeqChunk];
    float acc = 0.0f;
    const long long chunks = (npx + kSeqChunk - 1) / kSeqChunk;
    for (int k = threadIdx.x; k < kSeqChunk; k += blockDim.x) buf[0][k] = k < npx ? __ldcs(disp + k) : 0.0f;
    __syncthreads();
    for (long long c = 0; c < chunks; ++c) {
        const int cur = (int)(c & 1);
        const long long next0 = (c + 1) * kSeqChunk;
        if (threadIdx.x == 0) {
            const int n = (int)min((long long)kSeqChunk, npx - c * kSeqChunk);
            const float *b = buf[cur];
#pragma unroll 8
            for (int k = 0; k < n; ++k) acc += b[k];
        } else if (c + 1 < chunks) {
            for (int k = threadIdx.x - 1; k < kSeqChunk; k += blockDim.x - 1)
                buf[cur ^ 1][k] = next0 + k < npx ? __ldcs(disp + next0 + k) : 0.0f;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = acc;
}

// GaussBlur, main.cpp:11-33: window radius rs, clamp-to-edge, weights by squared distance from the host-built table
// (wtab[dsq], dsq <= 2 rs^2), accumulation order iy outer / ix inner, result round(val / wsum) per channel.
__global__ void __launch_bounds__(256)
k_gauss(const float *__restrict__ src3, int W, int H, int rs, const float *__restrict__ wtab, float *__restrict__ dst3) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;  // x
    const int i = blockIdx.y;                             // y
    if (j >= W) return;
    float ax = 0.0f, ay = 0.0f, az = 0.0f, wsum = 0.0f;
    for (int iy = i - rs; iy <= i + rs; ++iy) {
        const int y = min(H - 1, max(0, iy));
        for (int ix = j - rs; ix <= j + rs; ++ix) {
            const int x = min(W - 1, max(0, ix));
            const float w = __ldg(wtab + (ix - j) * (ix - j) + (iy - i) * (iy - i));
            const float *c = src3 + 3 * ((size_t)y * W + x);
            ax = ax + c[0] * w; ay = ay + c[1] * w; az = az + c[2] * w;
            wsum += w;
        }
    }
    float *o = dst3 + 3 * ((size_t)i * W + j);
    o[0] = roundf(ax / wsum); o[1] = roundf(ay / wsum); o[2] = roundf(az / wsum);
}

// MedianFilter, main.cpp:49-80: per channel, element w*w/2 (0-based) of the ascending (2w+1)^2 clamp-to-edge window —
// not the true median; kept.  An order statistic needs no particular visiting order: each thread keeps the k+1
// smallest values seen so far (insertion into a sorted prefix).
constexpr int kMedianMaxW = 15, kMedianMaxK = kMedianMaxW * kMedianMaxW / 2 + 1;
__global__ void __launch_bounds__(128)
k_median(const float *__restrict__ src3, int W, int H, int w, float *__restrict__ dst3) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, ch = blockIdx.z;
    if (x >= W) return;
    const int k = w * w / 2;  // wanted rank
    float best[kMedianMaxK];
    int have = 0;
    for (int dx = -w; dx <= w; ++dx) {
        const int i = max(min(dx + x, W - 1), 0);
        for (int dy = -w; dy <= w; ++dy) {
            const int jy = max(min(dy + y, H - 1), 0);
            const float v = src3[3 * ((size_t)jy * W + i) + ch];
            if (have <= k) {  // fill
                int p = have++;
                while (p > 0 && best[p - 1] > v) { best[p] = best[p - 1]; --p; }
                best[p] = v;
            } else if (v < best[k]) {
                int p = k;
                while (p > 0 && best[p - 1] > v) { best[p] = best[p - 1]; --p; }
                best[p] = v;
            }
        }
    }
    dst3[3 * ((size_t)y * W + x) + ch] = best[k];
}

// bitmap_image::set_pixel(x, y, r, g, b) with float arguments (main.cpp:193-201): the C++ float -> unsigned char
// conversion as x86-64 performs it (cvttss2si, low byte; out-of-range and NaN give 0), only where the pixel has samples.
__device__ __forceinline__ unsigned char to_u8(float v) {
    if (!(fabsf(v) < 2147483648.0f)) return 0;
    return (unsigned char)((int)v & 0xff);
}
__global__ void k_to_rgb8(const float *__restrict__ color3, const int *__restrict__ count, long long npx,
                          unsigned char *__restrict__ rgb8) {
    const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (p >= npx) return;
    const bool on = count[p] != 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) rgb8[3 * p + c] = on ? to_u8(color3[3 * p + c]) : 0;
}

}  // namespace ptd
