// SimCLR's two-view augmentation on the GPU.
//
// Replaces, for a batch of uint8 HWC images already on the device, what the reference's data workers do per
// image on the host with PIL (dataset.py:19-50, SimCLRTransforms -> create_simclr_data_augmentation):
//   RandomResizedCrop(size)  ->  RandomHorizontalFlip(0.5)  ->  RandomApply([ColorJitter(0.8 s, 0.8 s, 0.8 s, 0.2 s)], 0.8)
//   ->  RandomGrayscale(0.2)  ->  ToTensor,                              twice per image (two views).
// One CTA per (image, view).  The random parameters are drawn on the device by one thread from Philox4x32-10
// keyed by (seed, image, view, call offset) with torchvision's sampling rules (RandomResizedCrop.get_params with
// its ten attempts and centre-crop fallback, ColorJitter.get_params with a random order of the four adjustments),
// and are optionally written out, so a test can replay the same parameters through torchvision itself.
// The pixel arithmetic follows PIL's 8-bit paths step by step, rounding included: the two-pass antialiased bilinear
// resize of Resample.c (22-bit fixed-point coefficients, 8-bit intermediate), Image.blend's truncation for
// brightness / contrast / saturation, the RGB <-> HSV conversions of Convert.c for the hue shift, the ITU-R 601-2
// luma of convert("L").  Every stage works on uint8 pixels in shared memory; only ToTensor's division by 255
// produces floats.  Images up to 64 x 64 (CIFAR is 32 x 32), output size up to 64.
#include "ntx_common.cuh"

namespace {

constexpr int kThreads = 128;
constexpr int kMaxSide = 64;
constexpr int kParams = 16;     // floats written per (image, view) when the caller asks for the parameters

// ---- Philox4x32-10
struct Philox {
    uint32_t key[2], ctr[4], out[4];
    int have;
    __device__ void init(uint64_t seed, uint64_t a, uint64_t b) {
        key[0] = (uint32_t)seed; key[1] = (uint32_t)(seed >> 32);
        ctr[0] = (uint32_t)a; ctr[1] = (uint32_t)(a >> 32); ctr[2] = (uint32_t)b; ctr[3] = (uint32_t)(b >> 32);
        have = 0;
    }
    __device__ void round10() {
        uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
        if (++ctr[0] == 0) ++ctr[1];
        have = 4;
    }
    __device__ float uniform() {          // [0, 1)
        if (!have) round10();
        return (float)(out[--have] >> 8) * (1.0f / 16777216.0f);
    }
    __device__ int randint(int n) {       // [0, n)
        const int v = (int)(uniform() * (float)n);
        return v < n ? v : n - 1;
    }
};

struct Aug {
    int top, left, h, w;        // crop box
    int flip, jitter, gray;
    float bright, contrast, sat, hue;
    int order[4];               // 0 brightness, 1 contrast, 2 saturation, 3 hue
};

// torchvision RandomResizedCrop.get_params(scale = (0.08, 1), ratio = (3/4, 4/3)) and the rest of the pipeline's draws
__device__ void sample(Philox& rng, int H, int W, float strength, Aug& a) {
    const float area = (float)(H * W);
    const float log_lo = logf(3.f / 4.f), log_hi = logf(4.f / 3.f);
    bool ok = false;
    for (int t = 0; t < 10 && !ok; ++t) {
        const float target = area * (0.08f + rng.uniform() * (1.0f - 0.08f));
        const float aspect = expf(log_lo + rng.uniform() * (log_hi - log_lo));
        const int w = (int)rintf(sqrtf(target * aspect)), h = (int)rintf(sqrtf(target / aspect));
        if (w > 0 && w <= W && h > 0 && h <= H) {
            a.top = rng.randint(H - h + 1);
            a.left = rng.randint(W - w + 1);
            a.h = h; a.w = w;
            ok = true;
        }
    }
    if (!ok) {                  // centre crop with the ratio clamped
        const float in_ratio = (float)W / (float)H;
        int w = W, h = H;
        if (in_ratio < 3.f / 4.f) h = (int)rintf((float)w / (3.f / 4.f));
        else if (in_ratio > 4.f / 3.f) w = (int)rintf((float)h * (4.f / 3.f));
        a.h = h; a.w = w; a.top = (H - h) / 2; a.left = (W - w) / 2;
    }
    a.flip = rng.uniform() < 0.5f;
    a.jitter = !(0.8f < rng.uniform());                       // RandomApply: skipped when p < rand
    const float b = 0.8f * strength, hh = 0.2f * strength;
    for (int i = 0; i < 4; ++i) a.order[i] = i;
    for (int i = 3; i > 0; --i) {                             // random order of the four adjustments
        const int j = rng.randint(i + 1), t = a.order[i];
        a.order[i] = a.order[j];
        a.order[j] = t;
    }
    const float lo = fmaxf(0.f, 1.f - b);
    a.bright = lo + rng.uniform() * (1.f + b - lo);
    a.contrast = lo + rng.uniform() * (1.f + b - lo);
    a.sat = lo + rng.uniform() * (1.f + b - lo);
    a.hue = -hh + rng.uniform() * 2.f * hh;
    a.gray = rng.uniform() < 0.2f;
}

__device__ __forceinline__ uint8_t luma(uint8_t r, uint8_t g, uint8_t b) {           // PIL convert("L")
    return (uint8_t)((r * 19595u + g * 38470u + b * 7471u + 0x8000u) >> 16);
}
// PIL Image.blend(degenerate, image, alpha) on one channel (Blend.c): truncation, clipping only when extrapolating
__device__ __forceinline__ uint8_t blend(uint8_t degenerate, uint8_t px, float alpha) {
    const float t = (float)degenerate + alpha * ((float)px - (float)degenerate);
    if (alpha >= 0.f && alpha <= 1.f) return (uint8_t)t;
    return t <= 0.f ? 0 : (t >= 255.f ? 255 : (uint8_t)t);
}
__device__ __forceinline__ uint8_t clip8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : (uint8_t)v); }

// Convert.c rgb2hsv / hsv2rgb (8-bit), with the hue shifted in between as torchvision's adjust_hue does
__device__ void hue_shift(uint8_t& r, uint8_t& g, uint8_t& b, uint8_t shift) {
    const uint8_t maxc = max(r, max(g, b)), minc = min(r, min(g, b));
    uint8_t uh = 0, us = 0;
    const uint8_t uv = maxc;
    if (minc != maxc) {
        // float variables, double expressions: as the C source evaluates them
        const float cr = (float)(maxc - minc);
        const float s = cr / (float)maxc;
        const float rc = (float)(maxc - r) / cr, gc = (float)(maxc - g) / cr, bc = (float)(maxc - b) / cr;
        float h;
        if (r == maxc) h = bc - gc;
        else if (g == maxc) h = (float)(2.0 + rc - bc);
        else h = (float)(4.0 + gc - rc);
        h = (float)fmod((double)h / 6.0 + 1.0, 1.0);
        uh = clip8((int)((double)h * 255.0));
        us = clip8((int)((double)s * 255.0));
    }
    uh = (uint8_t)(uh + shift);                               // wraps, as the uint8 addition in adjust_hue does
    if (us == 0) {
        r = g = b = uv;
        return;
    }
    const int i = (int)floor((double)(float)uh * 6.0 / 255.0);
    const float f = (float)((double)(float)uh * 6.0 / 255.0 - (double)(float)i), fs = (float)((double)(float)us / 255.0);
    const double v = (double)(float)uv;
    const uint8_t p = clip8((int)round(v * (1.0 - (double)fs)));
    const uint8_t q = clip8((int)round(v * (1.0 - (double)fs * (double)f)));
    const uint8_t t = clip8((int)round(v * (1.0 - (double)fs * (1.0 - (double)f))));
    switch (i % 6) {
        case 0: r = uv; g = t; b = p; break;
        case 1: r = q; g = uv; b = p; break;
        case 2: r = p; g = uv; b = t; break;
        case 3: r = p; g = q; b = uv; break;
        case 4: r = t; g = p; b = uv; break;
        default: r = uv; g = p; b = q; break;
    }
}

// Resample.c: coefficients of output position xx for an input of in_size resampled to out_size (bilinear, antialiased)
__device__ void taps_for(int xx, int in_size, int out_size, int& xmin, int& n, int* k) {
    const double scale = (double)in_size / (double)out_size;
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = 1.0 * filterscale, ss = 1.0 / filterscale;
    const double center = (xx + 0.5) * scale;
    int lo = (int)(center - support + 0.5), hi = (int)(center + support + 0.5);
    if (lo < 0) lo = 0;
    if (hi > in_size) hi = in_size;
    xmin = lo;
    n = hi - lo;
    double ww = 0.0;
    for (int x = 0; x < n; ++x) {
        double t = (x + lo - center + 0.5) * ss;
        if (t < 0) t = -t;
        ww += t < 1.0 ? 1.0 - t : 0.0;
    }
    for (int x = 0; x < n; ++x) {
        double t = (x + lo - center + 0.5) * ss;
        if (t < 0) t = -t;
        double w = t < 1.0 ? 1.0 - t : 0.0;
        if (ww != 0.0) w /= ww;
        k[x] = (int)(0.5 + w * (double)(1 << 22));            // normalize_coeffs_8bpc (weights are non-negative)
    }
}

__global__ void __launch_bounds__(kThreads)
simclr_augment_kernel(const uint8_t* __restrict__ images, int batch, int H, int W, float strength, uint64_t seed,
                      uint64_t offset, int S, float* __restrict__ view0, float* __restrict__ view1,
                      float* __restrict__ params_out) {
    __shared__ uint8_t src[kMaxSide * kMaxSide * 3];          // the crop, then scratch
    __shared__ uint8_t mid[kMaxSide * kMaxSide * 3];          // after the horizontal pass: crop_h x S
    __shared__ uint8_t img[kMaxSide * kMaxSide * 3];          // S x S working image
    __shared__ Aug aug;
    __shared__ unsigned int lsum;
    const int image = blockIdx.x >> 1, view = blockIdx.x & 1, tid = threadIdx.x;
    if (tid == 0) {
        Philox rng;
        rng.init(seed, (uint64_t)image * 2 + view, offset);
        sample(rng, H, W, strength, aug);
        if (params_out) {
            float* p = params_out + (size_t)blockIdx.x * kParams;
            p[0] = (float)aug.top; p[1] = (float)aug.left; p[2] = (float)aug.h; p[3] = (float)aug.w;
            p[4] = (float)aug.flip; p[5] = (float)aug.jitter; p[6] = aug.bright; p[7] = aug.contrast;
            p[8] = aug.sat; p[9] = aug.hue; p[10] = (float)aug.gray;
            for (int i = 0; i < 4; ++i) p[11 + i] = (float)aug.order[i];
            p[15] = 0.f;
        }
    }
    __syncthreads();
    const Aug a = aug;
    // ---- crop into shared memory (h x w x 3)
    const uint8_t* base = images + (size_t)image * H * W * 3;
    for (int i = tid; i < a.h * a.w; i += kThreads) {
        const int y = i / a.w, x = i % a.w;
        const uint8_t* px = base + ((size_t)(a.top + y) * W + a.left + x) * 3;
        src[i * 3 + 0] = px[0]; src[i * 3 + 1] = px[1]; src[i * 3 + 2] = px[2];
    }
    __syncthreads();
    // ---- PIL resize((S, S), BILINEAR): horizontal pass (skipped when the width already matches), 8-bit result
    if (a.w != S) {
        for (int xx = tid; xx < S; xx += kThreads) {
            int xmin, n, k[2 * kMaxSide + 3];
            taps_for(xx, a.w, S, xmin, n, k);
            for (int y = 0; y < a.h; ++y)
                for (int c = 0; c < 3; ++c) {
                    int ss = 1 << 21;
                    for (int x = 0; x < n; ++x) ss += (int)src[(y * a.w + xmin + x) * 3 + c] * k[x];
                    mid[(y * S + xx) * 3 + c] = clip8(ss >> 22);
                }
        }
    } else {
        for (int i = tid; i < a.h * S * 3; i += kThreads) mid[i] = src[i];
    }
    __syncthreads();
    // ---- vertical pass
    if (a.h != S) {
        for (int yy = tid; yy < S; yy += kThreads) {
            int ymin, n, k[2 * kMaxSide + 3];
            taps_for(yy, a.h, S, ymin, n, k);
            for (int x = 0; x < S; ++x)
                for (int c = 0; c < 3; ++c) {
                    int ss = 1 << 21;
                    for (int y = 0; y < n; ++y) ss += (int)mid[((ymin + y) * S + x) * 3 + c] * k[y];
                    img[(yy * S + x) * 3 + c] = clip8(ss >> 22);
                }
        }
    } else {
        for (int i = tid; i < S * S * 3; i += kThreads) img[i] = mid[i];
    }
    __syncthreads();
    // ---- horizontal flip
    if (a.flip) {
        for (int i = tid; i < S * (S / 2); i += kThreads) {
            const int y = i / (S / 2), x = i % (S / 2);
            for (int c = 0; c < 3; ++c) {
                const uint8_t t = img[(y * S + x) * 3 + c];
                img[(y * S + x) * 3 + c] = img[(y * S + S - 1 - x) * 3 + c];
                img[(y * S + S - 1 - x) * 3 + c] = t;
            }
        }
        __syncthreads();
    }
    // ---- colour jitter: the four adjustments in the sampled order, on 8-bit pixels like PIL
    if (a.jitter) {
        for (int step = 0; step < 4; ++step) {
            const int op = a.order[step];
            if (op == 1) {                                    // contrast: blend with the mean grey level
                if (tid == 0) lsum = 0;
                __syncthreads();
                unsigned int part = 0;
                for (int i = tid; i < S * S; i += kThreads) part += luma(img[i * 3], img[i * 3 + 1], img[i * 3 + 2]);
                atomicAdd(&lsum, part);
                __syncthreads();
                const uint8_t mean = (uint8_t)(int)((double)lsum / (double)(S * S) + 0.5);
                for (int i = tid; i < S * S * 3; i += kThreads) img[i] = blend(mean, img[i], a.contrast);
            } else if (op == 0) {                             // brightness: blend with black
                for (int i = tid; i < S * S * 3; i += kThreads) img[i] = blend(0, img[i], a.bright);
            } else if (op == 2) {                             // saturation: blend with the pixel's own grey level
                for (int i = tid; i < S * S; i += kThreads) {
                    const uint8_t l = luma(img[i * 3], img[i * 3 + 1], img[i * 3 + 2]);
                    for (int c = 0; c < 3; ++c) img[i * 3 + c] = blend(l, img[i * 3 + c], a.sat);
                }
            } else {                                          // hue: shift H of the 8-bit HSV image
                const uint8_t shift = (uint8_t)((int)(a.hue * 255.0f) & 0xFF);
                for (int i = tid; i < S * S; i += kThreads) hue_shift(img[i * 3], img[i * 3 + 1], img[i * 3 + 2], shift);
            }
            __syncthreads();
        }
    }
    // ---- random grayscale (three equal channels), ToTensor: CHW float in [0, 1]
    float* out = (view == 0 ? view0 : view1) + (size_t)image * 3 * S * S;
    for (int i = tid; i < S * S; i += kThreads) {
        uint8_t r = img[i * 3], g = img[i * 3 + 1], b = img[i * 3 + 2];
        if (a.gray) r = g = b = luma(r, g, b);
        out[i] = (float)r / 255.0f;
        out[S * S + i] = (float)g / 255.0f;
        out[2 * S * S + i] = (float)b / 255.0f;
    }
}

}  // namespace

extern "C" int ntx_simclr_augment(const void* images, int batch, int height, int width, float strength,
                                  unsigned long long seed, unsigned long long offset, int out_size, float* view0,
                                  float* view1, float* params_out, void* stream) {
    NTX_CHECK_ARG(images && view0 && view1, "NULL buffer");
    NTX_CHECK_ARG(batch >= 1, "batch=%d must be positive", batch);
    NTX_CHECK_ARG(height >= 1 && height <= kMaxSide && width >= 1 && width <= kMaxSide,
                  "image size %dx%d outside [1, %d] (the image is staged in shared memory)", height, width, kMaxSide);
    NTX_CHECK_ARG(out_size >= 2 && out_size <= kMaxSide, "out_size=%d outside [2, %d]", out_size, kMaxSide);
    NTX_CHECK_ARG(strength > 0.f, "strength must be positive (main.py:49)");
    ntx_note_launch();
    simclr_augment_kernel<<<2 * batch, kThreads, 0, (cudaStream_t)stream>>>(
        static_cast<const uint8_t*>(images), batch, height, width, strength, (uint64_t)seed, (uint64_t)offset, out_size,
        view0, view1, params_out);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}
