// api.cu -- the C ABI of include/sr459.h: plans, stage scheduling, stand-alone operators.
#include "plan.h"

#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <new>
#include <string>

namespace sr459 {

std::atomic<uint64_t> g_launches{0};
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// numpy's pairwise summation (np.sum / np.mean on a contiguous 1-d float64 run), so that PSF
// normalisation reproduces the reference's rounding.
double np_pairwise_sum(const double* a, size_t n) {
    if (n < 8) {
        double r = 0.0;
        for (size_t i = 0; i < n; ++i) r += a[i];
        return r;
    }
    if (n <= 128) {
        double r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        size_t i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    size_t n2 = n / 2;
    n2 -= n2 % 8;
    return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
}

}  // namespace sr459

using namespace sr459;

// =================================================================================================
// plan
// =================================================================================================
int plan_ensure_partial(sr459_plan* p, size_t count) {
    if (count <= p->partial_cap) return SR459_OK;
    if (p->d_partial) { gfree(p->d_partial); p->scratch -= (int64_t)(p->partial_cap * sizeof(double)); p->d_partial = nullptr; }
    SR_TRY(plan_alloc(p, &p->d_partial, count));
    p->partial_cap = count;
    return SR459_OK;
}

int plan_ensure_errors(sr459_plan* p, int n_iter) {
    if (n_iter <= p->errors_cap) return SR459_OK;
    if (p->d_errors) { gfree(p->d_errors); p->scratch -= (int64_t)(p->errors_cap * sizeof(double)); p->d_errors = nullptr; }
    SR_TRY(plan_alloc(p, &p->d_errors, (size_t)n_iter));
    p->errors_cap = n_iter;
    return SR459_OK;
}

int plan_args_check(int n_frames, int h, int w, int factor, const double* shifts, const double* psf, int kh, int kw) {
    if (n_frames < 1 || n_frames > kMaxFrames) { set_error("n_frames must be 1..%d, got %d", kMaxFrames, n_frames); return SR459_EINVAL; }
    if (h < 1 || w < 1 || factor < 1) { set_error("bad geometry h=%d w=%d factor=%d", h, w, factor); return SR459_EINVAL; }
    if (!shifts || !psf || kh < 1 || kw < 1) { set_error("null shifts/psf or empty kernel"); return SR459_EINVAL; }
    return SR459_OK;
}

int plan_common_init(sr459_plan* p, int device, int n_frames, int h, int w, int factor,
                            const double* shifts, const double* psf, int kh, int kw, double step,
                            double lo, double hi) {
    SR_TRY(plan_args_check(n_frames, h, w, factor, shifts, psf, kh, kw));
    p->device = device; p->n = n_frames; p->h = h; p->w = w; p->factor = factor;
    p->H = h * factor; p->W = w * factor;
    p->psf.assign(psf, psf + (size_t)kh * kw); p->kh = kh; p->kw = kw;
    p->step = step; p->lo = lo; p->hi = hi;
    p->is_int = true; p->smax = 0;
    for (int k = 0; k < n_frames; ++k) {
        p->shifts[2 * k] = shifts[2 * k]; p->shifts[2 * k + 1] = shifts[2 * k + 1];
        double sy = shifts[2 * k] * factor, sx = shifts[2 * k + 1] * factor;
        if (sy != rint(sy) || sx != rint(sx) || fabs(sy) > 1e6 || fabs(sx) > 1e6) p->is_int = false;
        else {
            p->isy[k] = (int)rint(sy); p->isx[k] = (int)rint(sx);
            p->smax = std::max(p->smax, std::max(abs(p->isy[k]), abs(p->isx[k])));
        }
    }
    p->own_i0 = 0; p->own_i1 = h; p->hr_lo = 0; p->hr_hi = p->H; p->lr_lo = 0; p->lr_hi = h;

    build_taps_host(psf, kh, kw, false, p->taps_f);
    build_taps_host(psf, kh, kw, true, p->taps_b);
    SR_TRY(upload_taps(p->taps_f));
    SR_TRY(upload_taps(p->taps_b));
    if (kh == kw && (kh % 2) == 1 && kh <= 7) {
        int r = kh / 2;
        double k7[49] = {0};
        for (int u = 0; u < kh; ++u)
            for (int v = 0; v < kw; ++v) {
                k7[(u + 3 - r) * 7 + (v + 3 - r)] = psf[u * kw + v];
                p->psf7[(u + 3 - r) * 7 + (v + 3 - r)] = (float)psf[u * kw + v];
            }
        // rank-1 test in double: psf == column(through the peak) x row(through the peak) / peak
        int pu = 0, pv = 0;
        for (int i = 0; i < 49; ++i) if (fabs(k7[i]) > fabs(k7[pu * 7 + pv])) { pu = i / 7; pv = i % 7; }
        const double peak = k7[pu * 7 + pv];
        if (peak != 0.0) {
            double worst = 0.0;
            for (int u = 0; u < 7; ++u)
                for (int v = 0; v < 7; ++v)
                    worst = std::max(worst, fabs(k7[u * 7 + v] - k7[u * 7 + pv] * k7[pu * 7 + v] / peak));
            if (worst <= 1e-9 * fabs(peak)) {
                p->separable = true;
                for (int u = 0; u < 7; ++u) { p->acol[u] = (float)k7[u * 7 + pv]; p->brow[u] = (float)(k7[pu * 7 + u] / peak); }
            }
        }
    }
    if (fused_supported(p) && lat_supported(kh, kw, p->smax, p->W)) {
        SR_TRY(lat_create(&p->lat, psf, kh, kw));
        if (!p->side) {                                    // k_lat_edge_b runs beside k_lat_fwd
            SR_CUDA(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
            SR_CUDA(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
            SR_CUDA(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
        }
    }
    // parity lattices of the integer HR shifts (only meaningful for |shift| <= 1)
    if (p->is_int && p->smax <= 1 && factor == 2) {
        for (int ph = 0; ph < 4; ++ph) {
            int cnt = 0, ks[kMaxFrames];
            for (int k = 0; k < n_frames; ++k)
                if ((((p->isy[k] & 1) << 1) | (p->isx[k] & 1)) == ph) ks[cnt++] = k;
            if (!cnt) continue;
            const int s = p->nslot++;
            p->slot_pq[s] = ph; p->slot_n[s] = cnt;
            for (int f = 0; f < cnt; ++f) p->slot_k[s][f] = ks[f];
            // frames that still have a sample on the last site row / column: those with offset 0
            for (int ra = 0; ra < 2; ++ra)
                for (int cb = 0; cb < 2; ++cb) {
                    int m = 0;
                    for (int f = 0; f < cnt; ++f) {
                        const int oy = ((ph >> 1) + p->isy[ks[f]]) >> 1, ox = ((ph & 1) + p->isx[ks[f]]) >> 1;
                        if ((!ra || oy == 0) && (!cb || ox == 0)) ++m;
                    }
                    p->mtab[s][ra * 2 + cb] = (float)m;
                }
        }
    }
    return SR459_OK;
}

void plan_free(sr459_plan* p) {
    if (!p) return;
    DeviceGuard guard(p->device);
    void* ptrs[] = {p->d_zoom_y, p->d_zoom_x, p->d_saa_ty, p->d_saa_tx, p->d_fwd_ty, p->d_fwd_tx, p->d_bwd_ty,
                    p->d_bwd_tx, p->d_coef_lr, p->d_zoomed, p->d_coef_pad, p->d_B, p->d_acc, p->d_err, p->d_hr2,
                    p->d_partial, p->d_errors, p->d_edge, p->d_fir_iy0, p->d_fir_ix0, p->d_fir_wy, p->d_fir_wx, p->d_s1[0], p->d_s1[1], p->d_s1[2], p->d_s1[3], p->d_vsum,
                    p->d_edge_ml[0], p->d_edge_ml[1], p->d_edge_ml[2], p->d_edge_ml[3]};
    for (void* q : ptrs) if (q) gfree(q);
    free_taps(p->taps_f);
    free_taps(p->taps_b);
    lat_destroy(p->lat);
    band_state_free(p->bs);
    if (p->ev_fork) cudaEventDestroy(p->ev_fork);
    if (p->ev_join) cudaEventDestroy(p->ev_join);
    if (p->side) cudaStreamDestroy(p->side);
    delete p;
}

// ---- tables -------------------------------------------------------------------------------------
static int plan_zoom_tables(sr459_plan* p) {
    if (p->d_zoom_y) return SR459_OK;
    std::vector<AxisTap> ty, tx;
    build_zoom_axis(p->h, p->H, ty);
    build_zoom_axis(p->w, p->W, tx);
    SR_TRY(upload_taps_tbl(ty, &p->d_zoom_y));
    SR_TRY(upload_taps_tbl(tx, &p->d_zoom_x));
    return SR459_OK;
}

static int plan_saa_tables(sr459_plan* p) {
    if (p->d_saa_ty) return SR459_OK;
    std::vector<AxisTap> ty((size_t)p->n * p->H), tx((size_t)p->n * p->W), one;
    if (p->is_int) {
        // integer HR shift: shift(zoom(lr), s)[y] = zoom(lr)[clamp(y - s)]  -> re-indexed zoom taps
        std::vector<AxisTap> zy, zx;
        build_zoom_axis(p->h, p->H, zy);
        build_zoom_axis(p->w, p->W, zx);
        for (int k = 0; k < p->n; ++k) {
            for (int y = 0; y < p->H; ++y) ty[(size_t)k * p->H + y] = zy[std::min(std::max(y - p->isy[k], 0), p->H - 1)];
            for (int x = 0; x < p->W; ++x) tx[(size_t)k * p->W + x] = zx[std::min(std::max(x - p->isx[k], 0), p->W - 1)];
        }
    } else {
        for (int k = 0; k < p->n; ++k) {
            build_shift_axis(p->H, p->shifts[2 * k] * p->factor, p->H, 1, one);
            std::copy(one.begin(), one.end(), ty.begin() + (size_t)k * p->H);
            build_shift_axis(p->W, p->shifts[2 * k + 1] * p->factor, p->W, 1, one);
            std::copy(one.begin(), one.end(), tx.begin() + (size_t)k * p->W);
        }
    }
    SR_TRY(upload_taps_tbl(ty, &p->d_saa_ty));
    SR_TRY(upload_taps_tbl(tx, &p->d_saa_tx));
    return SR459_OK;
}

static int plan_generic_tables(sr459_plan* p) {
    if (p->d_fwd_ty) return SR459_OK;
    const int f = p->factor;
    std::vector<AxisTap> fy((size_t)p->n * p->h), fx((size_t)p->n * p->w), by((size_t)p->n * p->H), bx((size_t)p->n * p->W), one;
    for (int k = 0; k < p->n; ++k) {
        double sy = p->shifts[2 * k] * f, sx = p->shifts[2 * k + 1] * f;
        build_shift_axis(p->H, sy, p->h, f, one);  std::copy(one.begin(), one.end(), fy.begin() + (size_t)k * p->h);
        build_shift_axis(p->W, sx, p->w, f, one);  std::copy(one.begin(), one.end(), fx.begin() + (size_t)k * p->w);
        build_shift_axis(p->H, -sy, p->H, 1, one); std::copy(one.begin(), one.end(), by.begin() + (size_t)k * p->H);
        build_shift_axis(p->W, -sx, p->W, 1, one); std::copy(one.begin(), one.end(), bx.begin() + (size_t)k * p->W);
    }
    SR_TRY(upload_taps_tbl(fy, &p->d_fwd_ty));
    SR_TRY(upload_taps_tbl(fx, &p->d_fwd_tx));
    SR_TRY(upload_taps_tbl(by, &p->d_bwd_ty));
    SR_TRY(upload_taps_tbl(bx, &p->d_bwd_tx));
    // FIR tables of the back-projection (fir.cu): every frame, both axes; refused -> the spline stages stay in use
    if (f >= 2 && p->H == p->h * f && p->W == p->w * f) {
        std::vector<int> iy((size_t)p->n * p->H), ix((size_t)p->n * p->W), i1;
        std::vector<float> wy((size_t)p->n * p->H * kFirTaps), wx((size_t)p->n * p->W * kFirTaps), w1;
        bool ok = true;
        for (int k = 0; k < p->n && ok; ++k) {
            ok = build_fir_bp_axis(p->h, f, -p->shifts[2 * k] * f, i1, w1);
            if (ok) { std::copy(i1.begin(), i1.end(), iy.begin() + (size_t)k * p->H); std::copy(w1.begin(), w1.end(), wy.begin() + (size_t)k * p->H * kFirTaps); }
            ok = ok && build_fir_bp_axis(p->w, f, -p->shifts[2 * k + 1] * f, i1, w1);
            if (ok) { std::copy(i1.begin(), i1.end(), ix.begin() + (size_t)k * p->W); std::copy(w1.begin(), w1.end(), wx.begin() + (size_t)k * p->W * kFirTaps); }
        }
        if (ok) {
            SR_TRY(dev_alloc(&p->d_fir_iy0, iy.size())); SR_TRY(dev_alloc(&p->d_fir_ix0, ix.size()));
            SR_TRY(dev_alloc(&p->d_fir_wy, wy.size()));  SR_TRY(dev_alloc(&p->d_fir_wx, wx.size()));
            SR_CUDA(cudaMemcpy(p->d_fir_iy0, iy.data(), iy.size() * sizeof(int), cudaMemcpyHostToDevice));
            SR_CUDA(cudaMemcpy(p->d_fir_ix0, ix.data(), ix.size() * sizeof(int), cudaMemcpyHostToDevice));
            SR_CUDA(cudaMemcpy(p->d_fir_wy, wy.data(), wy.size() * sizeof(float), cudaMemcpyHostToDevice));
            SR_CUDA(cudaMemcpy(p->d_fir_wx, wx.data(), wx.size() * sizeof(float), cudaMemcpyHostToDevice));
            p->fir_ok = true;
            // SR459_NO_FIR_POLY=1 (read when the plan is built) keeps every tile on k_fir_bp: a debugging switch
            p->fir_poly_ok = (f == 2) && !getenv("SR459_NO_FIR_POLY") && fir_poly_analyse(p->n, p->H, p->W, iy, ix, wy, wx, p->fir_poly);
            if (p->fir_poly_ok) {
                SR_CUDA(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
                SR_CUDA(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
                SR_CUDA(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
            }
        }
    }
    return SR459_OK;
}

// ---- stage schedulers ------------------------------------------------------------------------------
static int run_zoom(sr459_plan* p, const float* d_lr, float* d_hr, cudaStream_t st) {
    SR_TRY(plan_zoom_tables(p));
    SR_TRY(plan_alloc(p, &p->d_coef_lr, (size_t)p->n * p->h * p->w));
    PrefilterJob j{d_lr, p->d_coef_lr, p->h, p->w, p->h, p->w, p->h, p->w, SRC_PLAIN, 1};
    SR_TRY(launch_prefilter2d(&j, 1, st));
    ResampleSumArgs a{};
    a.coef[0] = p->d_coef_lr; a.cpitch[0] = p->w; a.n = 1;
    a.ty = p->d_zoom_y; a.tx = p->d_zoom_x; a.Hout = p->H; a.Wout = p->W; a.scale = 1.0f; a.out = d_hr;
    return p->factor >= 2 ? launch_resample_tile(a, st) : launch_resample_sum(a, st);
}

static int run_saa(sr459_plan* p, const float* const* lr, float* d_hr, cudaStream_t st) {
    SR_TRY(plan_saa_tables(p));
    const size_t lpx = (size_t)p->h * p->w, hpx = (size_t)p->H * p->W;
    SR_TRY(plan_alloc(p, &p->d_coef_lr, (size_t)p->n * lpx));
    PrefilterJob jobs[kMaxFrames];
    for (int k = 0; k < p->n; ++k)
        jobs[k] = PrefilterJob{lr[k], p->d_coef_lr + k * lpx, p->h, p->w, p->h, p->w, p->h, p->w, SRC_PLAIN, 1};
    SR_TRY(launch_prefilter2d(jobs, p->n, st));
    ResampleSumArgs a{};
    a.n = p->n; a.Hout = p->H; a.Wout = p->W; a.out = d_hr; a.scale = 1.0f / (float)p->n;
    a.ty = p->d_saa_ty; a.tx = p->d_saa_tx;
    if (p->is_int) {
        for (int k = 0; k < p->n; ++k) { a.coef[k] = p->d_coef_lr + k * lpx; a.cpitch[k] = p->w; }
        return p->factor >= 2 ? launch_resample_tile(a, st) : launch_resample_sum(a, st);
    }
    // fractional: zoom every frame, pad+prefilter the zoomed frames, evaluate the shifted splines
    SR_TRY(plan_zoom_tables(p));
    const int PH = p->H + 2 * kPad, PW = p->W + 2 * kPad;
    SR_TRY(plan_alloc(p, &p->d_zoomed, (size_t)p->n * hpx));
    SR_TRY(plan_alloc(p, &p->d_coef_pad, (size_t)p->n * PH * PW));
    for (int k = 0; k < p->n; ++k) {
        ResampleSumArgs z{};
        z.coef[0] = p->d_coef_lr + k * lpx; z.cpitch[0] = p->w; z.n = 1;
        z.ty = p->d_zoom_y; z.tx = p->d_zoom_x; z.Hout = p->H; z.Wout = p->W; z.scale = 1.0f;
        z.out = p->d_zoomed + k * hpx;
        SR_TRY(launch_resample_sum(z, st));
        jobs[k] = PrefilterJob{p->d_zoomed + k * hpx, p->d_coef_pad + (size_t)k * PH * PW, p->H, p->W, p->H, p->W, PH, PW, SRC_EDGEPAD, 1};
    }
    SR_TRY(launch_prefilter2d(jobs, p->n, st));
    for (int k = 0; k < p->n; ++k) { a.coef[k] = p->d_coef_pad + (size_t)k * PH * PW; a.cpitch[k] = PW; }
    return launch_resample_sum(a, st);
}

// generic (stage-by-stage) IBP iteration: any shifts, any PSF, any integer factor
int iterate_generic(sr459_plan* p, const float* const* lr, const float* cur, float* nxt, int it, cudaStream_t st) {
    const size_t lpx = (size_t)p->h * p->w, hpx = (size_t)p->H * p->W;
    const int PH = p->H + 2 * kPad, PW = p->W + 2 * kPad;
    SR_TRY(plan_generic_tables(p));
    SR_TRY(plan_alloc(p, &p->d_B, hpx));
    SR_TRY(plan_alloc(p, &p->d_acc, hpx));
    SR_TRY(plan_alloc(p, &p->d_err, (size_t)p->n * lpx));
    SR_TRY(plan_alloc(p, &p->d_coef_pad, (size_t)p->n * PH * PW));
    // forward: blur, pad + prefilter once (shared by all frames), per-frame decimated evaluation
    CorrArgs c{cur, p->H, p->W, &p->taps_f, p->d_B, 0, nullptr, 0.f, 0.f, 0.f};
    SR_TRY(launch_corr2d(c, st));
    PrefilterJob jb{p->d_B, p->d_coef_pad, p->H, p->W, p->H, p->W, PH, PW, SRC_EDGEPAD, 1};
    SR_TRY(launch_prefilter2d(&jb, 1, st));
    ResampleErrArgs e{};
    e.coef = p->d_coef_pad; e.cpitch = PW; e.n = p->n; e.ty = p->d_fwd_ty; e.tx = p->d_fwd_tx; e.h = p->h; e.w = p->w;
    for (int k = 0; k < p->n; ++k) { e.lr[k] = lr[k]; e.out[k] = p->d_err + k * lpx; }
    e.partial = p->d_partial + (size_t)it * p->per_iter;
    e.mse_i0 = p->mse_i0 < 0 ? 0 : p->mse_i0; e.mse_i1 = p->mse_i1 < 0 ? p->h : p->mse_i1;
    SR_TRY(launch_resample_err_tile(e, st));
    // backward: zero-insert + pad + prefilter per frame, evaluate at -s, sum, transposed blur + update
    if (p->fir_ok && p->path == 0) {                        // the same linear map as ONE table-driven FIR kernel (fir.cu)
        FirBpArgs fb{};
        for (int k = 0; k < p->n; ++k) fb.err[k] = p->d_err + k * lpx;
        fb.n = p->n; fb.h = p->h; fb.w = p->w; fb.H = p->H; fb.W = p->W;
        fb.iy0 = p->d_fir_iy0; fb.ix0 = p->d_fir_ix0; fb.wy = p->d_fir_wy; fb.wx = p->d_fir_wx; fb.out = p->d_acc;
        if (p->fir_poly_ok) {                               // interior tiles: polyphase register-window kernel
            FirPolyArgs fp = p->fir_poly;
            for (int k = 0; k < p->n; ++k) fp.err[k] = fb.err[k];
            fp.h = p->h; fp.w = p->w; fp.out = p->d_acc;
            fb.skip_y0 = fp.y0; fb.skip_y1 = fp.y1; fb.skip_x0 = fp.x0; fb.skip_x1 = fp.x1;
            // fork / join on the plan's side stream (capturable): the ~150 border CTAs of k_fir_bp are one latency-bound
            // wave on their own, the interior kernel runs beside them
            SR_CUDA(cudaEventRecord(p->ev_fork, st));
            SR_CUDA(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
            SR_TRY(launch_fir_bp(fb, p->side));
            SR_CUDA(cudaEventRecord(p->ev_join, p->side));
            SR_TRY(launch_fir_poly(fp, st));
            SR_CUDA(cudaStreamWaitEvent(st, p->ev_join, 0));
        } else {
            SR_TRY(launch_fir_bp(fb, st));
        }
        CorrArgs u2{p->d_acc, p->H, p->W, &p->taps_b, nxt, 1, cur, (float)(p->step / p->n), (float)p->lo, (float)p->hi};
        if (p->store_y0 >= 0) { u2.y0 = p->store_y0; u2.y1 = p->store_y1; }
        return launch_corr2d(u2, st);
    }
    PrefilterJob jobs[kMaxFrames];
    for (int k = 0; k < p->n; ++k)
        jobs[k] = PrefilterJob{p->d_err + k * lpx, p->d_coef_pad + (size_t)k * PH * PW, p->h, p->w, p->H, p->W, PH, PW, SRC_ZI_EDGEPAD, p->factor};
    SR_TRY(launch_prefilter2d(jobs, p->n, st));
    ResampleSumArgs s{};
    s.n = p->n; s.Hout = p->H; s.Wout = p->W; s.scale = 1.0f; s.out = p->d_acc; s.ty = p->d_bwd_ty; s.tx = p->d_bwd_tx;
    for (int k = 0; k < p->n; ++k) { s.coef[k] = p->d_coef_pad + (size_t)k * PH * PW; s.cpitch[k] = PW; }
    SR_TRY(launch_resample_sum(s, st));
    CorrArgs u{p->d_acc, p->H, p->W, &p->taps_b, nxt, 1, cur, (float)(p->step / p->n), (float)p->lo, (float)p->hi};
    if (p->store_y0 >= 0) { u.y0 = p->store_y0; u.y1 = p->store_y1; }
    return launch_corr2d(u, st);
}

int generic_per_iter(const sr459_plan* p) { return ((p->h + kRsTileH - 1) / kRsTileH) * ((p->w + kRsTileW - 1) / kRsTileW); }   // k_resample_err_tile: one slot per CTA

// ---- fused integer path ------------------------------------------------------------------------------
bool fused_supported(const sr459_plan* p) {
    return p->is_int && p->factor == 2 && p->smax <= kMaxIntShift && std::max(p->taps_f.reach, p->taps_b.reach) <= 16;
}

static void fill_geom(const sr459_plan* p, IbpGeom& g, int uy0, int uy1) {
    g.H = p->H; g.W = p->W; g.h = p->h; g.w = p->w; g.n = p->n;
    for (int k = 0; k < kMaxFrames; ++k) { g.sy[k] = k < p->n ? p->isy[k] : 0; g.sx[k] = k < p->n ? p->isx[k] : 0; }
    g.smax = p->smax;
    g.alpha = (float)(p->step / p->n); g.lo = (float)p->lo; g.hi = (float)p->hi;
    g.hr_lo = p->hr_lo; g.hr_hi = p->hr_hi; g.lr_lo = p->lr_lo; g.lr_hi = p->lr_hi;
    g.upd_y0 = uy0; g.upd_y1 = uy1; g.own_i0 = p->own_i0; g.own_i1 = p->own_i1;
}


static int dense_launch(const sr459_plan* p, const IbpGeom& g, const float* const* lr, const float* cur, float* nxt,
                        const DensePass& dp, double* partial, int* ntiles_out, cudaStream_t st) {
    DenseArgs a{};
    a.g = g;
    for (int k = 0; k < p->n; ++k) a.lr[k] = lr[k];
    a.hr_in = cur; a.hr_out = nxt;
    a.wf = p->taps_f.d_w; a.fdy = p->taps_f.d_dy; a.fdx = p->taps_f.d_dx; a.nf = p->taps_f.ntaps;
    a.wb = p->taps_b.d_w; a.bdy = p->taps_b.d_dy; a.bdx = p->taps_b.d_dx; a.nb = p->taps_b.ntaps;
    a.R = std::max(p->taps_f.reach, p->taps_b.reach);
    a.tile_h = dp.tile_h; a.tile_w = dp.tile_w;
    a.nrect = 0; a.tiles_before[0] = 0;
    for (int r = 0; r < dp.nrect; ++r) {
        const Rect& rc = dp.rects[r];
        if (rc.y1 <= rc.y0 || rc.x1 <= rc.x0) continue;
        int i = a.nrect++;
        a.rect[i] = rc;
        a.tiles_x[i] = (rc.x1 - rc.x0 + dp.tile_w - 1) / dp.tile_w;
        int ty = (rc.y1 - rc.y0 + dp.tile_h - 1) / dp.tile_h;
        a.tiles_before[i + 1] = a.tiles_before[i] + a.tiles_x[i] * ty;
    }
    a.partial = partial;
    *ntiles_out = a.tiles_before[a.nrect];
    if (a.nrect == 0) return SR459_OK;
    return launch_ibp_dense(a, st);
}

static void fill_lat_call(const sr459_plan* p, LatCall& c, int uy0, int uy1) {
    c.H = p->H; c.W = p->W; c.h = p->h; c.w = p->w; c.n = p->n;
    for (int k = 0; k < p->n; ++k) { c.sy[k] = p->isy[k]; c.sx[k] = p->isx[k]; }
    c.hr_lo = p->hr_lo; c.hr_hi = p->hr_hi; c.lr_lo = p->lr_lo; c.lr_hi = p->lr_hi;
    c.own_i0 = p->own_i0; c.own_i1 = p->own_i1; c.upd_y0 = uy0; c.upd_y1 = uy1;
    c.alpha = (float)(p->step / p->n); c.lo = (float)p->lo; c.hi = (float)p->hi;
}

// number of partial-sum slots one fused iteration writes, and the launch recipe

FusedRecipe fused_recipe(const sr459_plan* p, int uy0, int uy1) {
    FusedRecipe r{};
    const int W = p->W;
    int fth, ftw;
    ibp_fast_tile(&fth, &ftw);
    bool fast_ok = ibp_fast_supported(p->kh, p->kw, p->smax) && (W % 4 == 0) && p->path != 2 && p->path != 1 && p->path != 8;
    bool sep_ok = ibp_sep_supported(p->kh, p->kw, p->smax, p->separable, W) && (p->path == 0 || p->path == 4 || p->path == 7);
    bool big = (long long)(uy1 - uy0) * W >= 64 * 64;
    const bool lat_ok = p->lat != nullptr && (p->path == 0 || p->path == 8);
    // the streaming kernel treats the local buffer (a row band with its halo) as an image of its own: what the
    // artificial cuts get wrong stays inside the halo rows, which the launch does not store (segments cover uy0..uy1)
    const bool stream_ok = sep_ok && (p->path == 0 || p->path == 7) && (p->hr_lo % 2 == 0) && 2 * p->lr_lo == p->hr_lo &&
                           (!p->band || (uy0 - p->hr_lo >= 8 || uy0 == 0) && (p->hr_hi - uy1 >= 8 || uy1 == p->H)) &&
                           ibp_stream_supported(p->kh, p->kw, p->smax, p->separable, W, p->nslot);
    if (stream_ok && (big || p->path == 7)) {
        r.use_stream = true;
        int n_sm = p->n_sm;
        if (!n_sm) { cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, p->device); }
        // row bands: the rows a neighbour needs sit in short border segments that run first (ibp_stream.cu)
        const int brows = 8 * kStreamBorderNit - 6;
        if (p->band && p->nranks > 1 && uy1 - uy0 >= 4 * brows) {
            r.st_bnit = kStreamBorderNit; r.st_btop = p->rank > 0; r.st_bbot = p->rank < p->nranks - 1;
        }
        stream_geometry(uy1 - uy0 - (r.st_btop + r.st_bbot) * brows, W, n_sm, &r.st_nit, &r.st_seg, &r.st_nseg, &r.st_nstrip);
        r.n_fast = (r.st_nseg + r.st_btop + r.st_bbot) * r.st_nstrip;
    } else if (sep_ok && (p->path == 0 || p->path == 7) && !p->band && uy0 == 0 && uy1 == p->H && (big || p->path == 7) &&
               ibp_stream_ml_supported(p->kh, p->kw, p->smax, p->separable, W, p->nslot)) {
        r.use_stream_ml = true;
        int n_sm = p->n_sm;
        if (!n_sm) { cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, p->device); }
        stream_ml_geometry(p->H, W, n_sm, &r.st_nit, &r.st_seg, &r.st_nseg, &r.st_nstrip);
        r.n_fast = r.st_nseg * r.st_nstrip;
    } else if (sep_ok && p->path != 7 && (big || p->path >= 4)) {
        r.use_ws = true;
        ibp_ws_tile(&fth, &ftw);
        r.use_sep = true;
        r.fast_tx = (W + ftw - 1) / ftw;
        r.fast_ty = (uy1 - uy0 + fth - 1) / fth;
        r.n_fast = r.fast_tx * r.fast_ty;
    } else if (lat_ok && (p->path == 8 || (p->path == 0 && !fast_ok && big))) {
        r.use_lat = true;
        LatCall c{}; fill_lat_call(p, c, uy0, uy1);
        r.n_fast = lat_per_iter(p->lat, c);
    } else if (fast_ok && p->path < 4 && (big || p->path == 3)) {
        r.use_fast = true;
        r.fast_tx = (W + ftw - 1) / ftw;
        r.fast_ty = (uy1 - uy0 + fth - 1) / fth;
        r.n_fast = r.fast_tx * r.fast_ty;
    } else {
        r.all.tile_h = 32; r.all.tile_w = 32; r.all.nrect = 1;
        r.all.rects[0] = Rect{uy0, 0, uy1, W};
        r.n_all = ((W + 31) / 32) * ((uy1 - uy0 + 31) / 32);
    }
    r.per_iter = r.n_fast + r.n_all;
    return r;
}

// S1 planes + sum of V for the frames of this call (ibp_sep.cu restructuring (2))
int sep_prepare(sr459_plan* p, const float* const* lr, bool stream_kernel, cudaStream_t st, bool stream_ml) {
    const int rows = p->lr_hi - p->lr_lo;
    const int pitch = (p->w + 3) & ~3;
    if (p->s1_rows != rows || p->s1_pitch != pitch) {
        for (int s = 0; s < kMaxSlots; ++s) if (p->d_s1[s]) { gfree(p->d_s1[s]); p->d_s1[s] = nullptr; }
        p->s1_rows = rows; p->s1_pitch = pitch;
    }
    SR_TRY(plan_alloc(p, &p->d_vsum, 1));
    LatticeArgs a{};
    for (int k = 0; k < p->n; ++k) a.lr[k] = lr[k];
    a.h = p->h; a.w = p->w; a.lr_lo = p->lr_lo; a.lr_hi = p->lr_hi;
    a.nslot = p->nslot;
    for (int s = 0; s < p->nslot; ++s) {
        if (!p->d_s1[s]) {
            SR_TRY(plan_alloc(p, &p->d_s1[s], (size_t)rows * pitch));
            SR_TRY(ws_make_s1_map(&p->tm_s1w[s], p->d_s1[s], p->w, rows, pitch));
        }
        a.s1[s] = p->d_s1[s];
        a.slot_n[s] = p->slot_n[s];
        const int P = p->slot_pq[s] >> 1, Q = p->slot_pq[s] & 1;
        for (int f = 0; f < p->slot_n[s]; ++f) {
            const int k = p->slot_k[s][f];
            a.slot_k[s][f] = k;
            a.slot_oy[s][f] = (P + p->isy[k]) >> 1;
            a.slot_ox[s][f] = (Q + p->isx[k]) >> 1;
        }
    }
    a.rows = rows; a.row0 = p->lr_lo; a.pitch = pitch;
    a.own_a0 = p->own_i0; a.own_a1 = p->own_i1;
    a.vsum = p->d_vsum;
    if (!p->n_sm) SR_CUDA(cudaDeviceGetAttribute(&p->n_sm, cudaDevAttrMultiProcessorCount, p->device));
    SR_TRY(launch_lattice_sums(a, st));
    if (stream_kernel) {                              // the streaming kernel's map and edge vectors (+ their variances)
        SR_TRY(stream_make_s1_map(&p->tm_s1s, p->d_s1[0], p->w, rows, pitch));
        SR_TRY(plan_alloc(p, &p->d_edge, (size_t)2 * p->h + 2 * p->w + 4));
        const int h_loc = (p->hr_hi - p->hr_lo) / 2;     // the local image of a row band (whole image: h)
        SR_TRY(launch_stream_edge_sums(lr, p->n, h_loc, p->w, p->isy, p->isx, p->slot_pq[0] >> 1, p->slot_pq[0] & 1,
                                       p->d_edge, p->d_vsum, p->own_i0 - p->lr_lo, p->own_i1 - p->lr_lo, p->own_i0 == 0, st));
    }
    if (stream_ml) {                                  // per lattice: map + edge vectors of ITS frames (+ their variances)
        for (int s = 0; s < p->nslot; ++s) {
            SR_TRY(stream_make_s1_map(&p->tm_s1ml[s], p->d_s1[s], p->w, rows, pitch));
            SR_TRY(plan_alloc(p, &p->d_edge_ml[s], (size_t)2 * p->h + 2 * p->w + 4));
            const float* sub[kMaxFrames]; int sy[kMaxFrames], sx[kMaxFrames];
            for (int f = 0; f < p->slot_n[s]; ++f) { const int k = p->slot_k[s][f]; sub[f] = lr[k]; sy[f] = p->isy[k]; sx[f] = p->isx[k]; }
            SR_TRY(launch_stream_edge_sums(sub, p->slot_n[s], p->h, p->w, sy, sx, p->slot_pq[s] >> 1, p->slot_pq[s] & 1,
                                           p->d_edge_ml[s], p->d_vsum, 0, p->h, 1, st));
        }
    }
    return SR459_OK;
}

int iterate_fused(sr459_plan* p, const FusedRecipe& r, const float* const* lr, const float* cur, float* nxt,
                  double* partial, int uy0, int uy1, cudaStream_t st, int it) {
    IbpGeom g;
    fill_geom(p, g, uy0, uy1);
    if (r.use_stream) {
        StreamArgs f{};
        const int H_loc = p->hr_hi - p->hr_lo, h_loc = H_loc / 2;
        SR_TRY(stream_make_hr_map(&f.tm_hr, cur, p->W, H_loc));
        f.tm_s1 = p->tm_s1s;
        f.hr_out = nxt;
        f.H = H_loc; f.W = p->W; f.h = h_loc; f.w = p->w;
        f.y_base = uy0 - p->hr_lo; f.y_end = uy1 - p->hr_lo;
        f.nit = r.st_nit; f.seg_rows = r.st_seg; f.nseg = r.st_nseg; f.nstrip = r.st_nstrip;
        f.b_nit = r.st_bnit; f.b_top = r.st_btop; f.b_bot = r.st_bbot;
        f.alpha = g.alpha; f.lo = g.lo; f.hi = g.hi;
        memcpy(f.acol, p->acol, sizeof(f.acol));
        memcpy(f.brow, p->brow, sizeof(f.brow));
        memcpy(f.mtab, p->mtab[0], sizeof(f.mtab));
        const int P = p->slot_pq[0] >> 1, Q = p->slot_pq[0] & 1;
        f.has_left = (Q == 1);
        f.colS1 = p->d_edge; f.leftS1 = p->d_edge + h_loc; f.rowS1 = p->d_edge + 2 * h_loc;
        f.topS1 = p->d_edge + 2 * h_loc + p->w; f.scal = p->d_edge + 2 * h_loc + 2 * p->w;
        for (int k = 0; k < p->n; ++k) {
            const int oy = (P + p->isy[k]) >> 1, ox = (Q + p->isx[k]) >> 1;
            if (p->isx[k] == -1) { f.mc[0] += 1.f; if (oy == 0) f.mc[1] += 1.f; }
            if (p->isx[k] == +1) { f.ml[0] += 1.f; if (oy == 0) f.ml[1] += 1.f; }
            if (p->isy[k] == -1) { f.mr[0] += 1.f; if (ox == 0) f.mr[1] += 1.f; }
            if (p->isy[k] == +1) { f.mt[0] += 1.f; if (ox == 0) f.mt[1] += 1.f; }
            if (p->isy[k] == -1 && p->isx[k] == -1) f.m_dupc += 1.f;
            if (p->isy[k] == +1 && p->isx[k] == +1) f.m_msec += 1.f;
        }
        f.partial = partial;
        if (p->kpush) f.bp = *p->kpush;                   // row bands: the kernel exchanges the halo rows itself
        return launch_ibp_stream(f, P, Q, st);
    }
    if (r.use_stream_ml) {
        StreamMLArgs f{};
        SR_TRY(stream_ml_make_hr_map(&f.tm_hr, cur, p->W, p->H));
        f.hr_out = nxt;
        f.H = p->H; f.W = p->W; f.h = p->h; f.w = p->w;
        f.nit = r.st_nit; f.seg_rows = r.st_seg; f.nseg = r.st_nseg; f.nstrip = r.st_nstrip;
        f.alpha = g.alpha; f.lo = g.lo; f.hi = g.hi;
        memcpy(f.acol, p->acol, sizeof(f.acol));
        memcpy(f.brow, p->brow, sizeof(f.brow));
        int mask = 0;
        for (int s = 0; s < p->nslot; ++s) {
            const int P = p->slot_pq[s] >> 1, Q = p->slot_pq[s] & 1;
            mask |= 1 << p->slot_pq[s];
            f.tm_s1[s] = p->tm_s1ml[s];
            memcpy(f.mtab[s], p->mtab[s], sizeof(f.mtab[s]));
            f.edge[s] = p->d_edge_ml[s];
            for (int i = 0; i < p->slot_n[s]; ++i) {
                const int k = p->slot_k[s][i];
                const int oy = (P + p->isy[k]) >> 1, ox = (Q + p->isx[k]) >> 1;
                if (p->isx[k] == -1) { f.mc[s][0] += 1.f; if (oy == 0) f.mc[s][1] += 1.f; }
                if (p->isx[k] == +1) { f.ml[s][0] += 1.f; if (oy == 0) f.ml[s][1] += 1.f; }
                if (p->isy[k] == -1) { f.mr[s][0] += 1.f; if (ox == 0) f.mr[s][1] += 1.f; }
                if (p->isy[k] == +1) { f.mt[s][0] += 1.f; if (ox == 0) f.mt[s][1] += 1.f; }
                if (p->isy[k] == -1 && p->isx[k] == -1) f.m_dupc[s] += 1.f;
                if (p->isy[k] == +1 && p->isx[k] == +1) f.m_msec[s] += 1.f;
            }
        }
        f.partial = partial;
        return launch_ibp_stream_ml(f, mask, st);
    }
    if (r.use_lat) {
        LatCall c{};
        fill_lat_call(p, c, uy0, uy1);
        for (int k = 0; k < p->n; ++k) c.lr[k] = lr[k];
        c.hr_in = cur; c.hr_out = nxt; c.partial = partial;
        if (p->lat_s1) {                                   // S1 planes of this call (sep_prepare): elementwise residual pass
            c.s1_n = p->nslot; c.s1_row0 = p->lr_lo; c.s1_rows = p->s1_rows;
            for (int s = 0; s < p->nslot; ++s) { c.s1_pq[s] = p->slot_pq[s]; c.s1[s] = p->d_s1[s]; }
        }
        c.side = p->side; c.ev_fork = p->ev_fork; c.ev_join = p->ev_join;
        return lat_iterate(p->lat, c, st);
    }
    if (r.use_sep) {
        SepArgs f{};
        SR_TRY(ws_make_hr_map(&f.tm_hr, cur, p->W, p->hr_hi - p->hr_lo));
        f.g = g;
        for (int k = 0; k < p->n; ++k) f.lr[k] = lr[k];
        f.hr_out = nxt;
        memcpy(f.psf, p->psf7, sizeof(f.psf));
        memcpy(f.acol, p->acol, sizeof(f.acol));
        memcpy(f.brow, p->brow, sizeof(f.brow));
        f.nph = p->nslot;
        for (int s = 0; s < p->nslot; ++s) {
            f.tm_s1[s] = p->tm_s1w[s];
            f.ph_pq[s] = p->slot_pq[s];
            memcpy(f.mtab[s], p->mtab[s], sizeof(f.mtab[s]));
        }
        f.s1_row0 = p->lr_lo;
        f.tiles_x = r.fast_tx; f.tiles_y = r.fast_ty; f.ty_origin = uy0;
        f.partial = partial;
        f.n_sm = p->n_sm;
        { static const int dbg = getenv("SR459_DBG") ? atoi(getenv("SR459_DBG")) : 0; f.dbg = dbg; }
        return launch_ibp_ws(f, st);
    }
    if (r.use_fast) {
        FastArgs f{};
        f.g = g;
        for (int k = 0; k < p->n; ++k) f.lr[k] = lr[k];
        f.hr_in = cur; f.hr_out = nxt;
        memcpy(f.psf, p->psf7, sizeof(f.psf));
        for (int k = 0; k < p->n; ++k) {
            int ph = ((p->isy[k] & 1) << 1) | (p->isx[k] & 1);
            f.ph_k[ph][f.ph_n[ph]++] = k;
        }
        f.tiles_x = r.fast_tx; f.tiles_y = r.fast_ty; f.ty_origin = uy0;
        f.partial = partial;
        return launch_ibp_fast(f, st);
    }
    int nt = 0;
    return dense_launch(p, g, lr, cur, nxt, r.all, partial, &nt, st);
}

bool use_fused(const sr459_plan* p) {
    if (p->path == 1) return false;
    return fused_supported(p);
}

static int run_ibp(sr459_plan* p, const float* const* lr, float* d_hr, int n_iter, cudaStream_t st) {
    if (n_iter < 0) { set_error("n_iter must be >= 0"); return SR459_EINVAL; }
    if (n_iter == 0) return SR459_OK;
    const size_t hpx = (size_t)p->H * p->W;
    SR_TRY(plan_alloc(p, &p->d_hr2, hpx));
    SR_TRY(plan_ensure_errors(p, n_iter));
    float* cur = d_hr; float* nxt = p->d_hr2;
    const double scale = 1.0 / ((double)p->n * p->h * p->w);
    const double* d_add = nullptr;
    if (use_fused(p)) {
        FusedRecipe r = fused_recipe(p, 0, p->H);
        p->per_iter = r.per_iter;
        SR_TRY(plan_ensure_partial(p, (size_t)n_iter * r.per_iter));
        p->lat_s1 = false;
        if (r.use_sep || r.use_stream || r.use_stream_ml || (r.use_lat && p->nslot > 0 && !getenv("SR459_LAT_NO_S1"))) {
            SR_TRY(sep_prepare(p, lr, r.use_stream, st, r.use_stream_ml)); d_add = p->d_vsum; p->lat_s1 = r.use_lat;
        }
        for (int it = 0; it < n_iter; ++it) {
            SR_TRY(iterate_fused(p, r, lr, cur, nxt, p->d_partial + (size_t)it * r.per_iter, 0, p->H, st, it));
            std::swap(cur, nxt);
        }
    } else {
        p->per_iter = generic_per_iter(p);
        SR_TRY(plan_ensure_partial(p, (size_t)n_iter * p->per_iter));
        for (int it = 0; it < n_iter; ++it) {
            SR_TRY(iterate_generic(p, lr, cur, nxt, it, st));
            std::swap(cur, nxt);
        }
    }
    SR_TRY(launch_reduce_partials(p->d_partial, p->per_iter, n_iter, scale, d_add, p->d_errors, st));
    if (cur != d_hr) SR_CUDA(cudaMemcpyAsync(d_hr, cur, hpx * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return SR459_OK;
}

int check_lr(const sr459_plan* p, const float* const* lr) {
    if (!p) { set_error("null plan"); return SR459_EINVAL; }
    if (!lr) { set_error("null frame pointer array"); return SR459_EINVAL; }
    for (int k = 0; k < p->n; ++k) if (!lr[k]) { set_error("frame %d pointer is null", k); return SR459_EINVAL; }
    return SR459_OK;
}

// =================================================================================================
// exported functions
// =================================================================================================
extern "C" {

#ifndef SR459_SOURCE_STAMP
#define SR459_SOURCE_STAMP "unstamped"
#endif
int sr459_version(void) { return 200; }
const char* sr459_source_stamp(void) { return SR459_SOURCE_STAMP; }
const char* sr459_last_error(void) { return g_err; }
uint64_t sr459_launch_count(void) { return g_launches.load(); }

int sr459_psf_gaussian(int size, double sigma, double* h_out) {
    if (size < 1 || !(sigma > 0.0) || !h_out) { set_error("psf_gaussian: bad arguments"); return SR459_EINVAL; }
    int hw = size / 2, K = 2 * hw + 1;
    for (int u = 0; u < K; ++u)
        for (int v = 0; v < K; ++v) {
            double y = u - hw, x = v - hw;
            h_out[u * K + v] = exp(-(x * x + y * y) / (2.0 * sigma * sigma));
        }
    double s = np_pairwise_sum(h_out, (size_t)K * K);
    for (int i = 0; i < K * K; ++i) h_out[i] /= s;
    return SR459_OK;
}

int sr459_psf_measured(const void* d_imgs, int is_f32, int n_img, int H, int W, int halfwidth, double* h_out,
                       int* h_peaks, int* h_n_used, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!d_imgs || n_img < 1 || H < 1 || W < 1 || halfwidth < 1 || !h_out) { set_error("psf_measured: bad arguments"); return SR459_EINVAL; }
    const int margin = halfwidth + 6, side = 2 * margin + 1, K = 2 * halfwidth + 1;
    long long* d_keys = nullptr; int* d_peaks = nullptr; double* d_patch = nullptr;
    SR_CUDA(gmalloc((void**)&d_keys, sizeof(long long) * n_img));
    SR_CUDA(gmalloc((void**)&d_peaks, sizeof(int) * 2 * n_img));
    SR_CUDA(gmalloc((void**)&d_patch, sizeof(double) * side * side));
    int rc = launch_argmax(d_imgs, is_f32, n_img, (int64_t)H * W, d_keys, st);
    std::vector<unsigned long long> keys(n_img);
    std::vector<int> peaks(2 * n_img, -1);
    std::vector<double> patch((size_t)side * side);
    int used = 0;
    cudaError_t ce = cudaSuccess;
    if (rc == SR459_OK) {
        ce = cudaMemcpyAsync(keys.data(), d_keys, sizeof(long long) * n_img, cudaMemcpyDeviceToHost, st);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
    }
    if (rc == SR459_OK && ce == cudaSuccess) {
        for (int k = 0; k < n_img; ++k) {
            unsigned idx = 0xffffffffu - (unsigned)(keys[k] & 0xffffffffULL);
            int pr = (int)(idx / (unsigned)W), pc = (int)(idx % (unsigned)W);
            // run_sr.py:128-131
            if (pr < margin || pr + margin + 1 > H || pc < margin || pc + margin + 1 > W) continue;
            peaks[2 * k] = pr; peaks[2 * k + 1] = pc; ++used;
        }
        ce = cudaMemcpyAsync(d_peaks, peaks.data(), sizeof(int) * 2 * n_img, cudaMemcpyHostToDevice, st);
        if (ce == cudaSuccess && used > 0) {
            rc = launch_patch_sum(d_imgs, is_f32, n_img, H, W, d_peaks, margin, d_patch, st);
            if (rc == SR459_OK) ce = cudaMemcpyAsync(patch.data(), d_patch, sizeof(double) * side * side, cudaMemcpyDeviceToHost, st);
        }
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
    }
    gfree(d_keys); gfree(d_peaks); gfree(d_patch);
    if (rc != SR459_OK) return rc;
    if (ce != cudaSuccess) { set_error("psf_measured: %s", cudaGetErrorString(ce)); return SR459_ECUDA; }
    if (h_peaks) memcpy(h_peaks, peaks.data(), sizeof(int) * 2 * n_img);
    if (h_n_used) *h_n_used = used;
    if (used == 0) { set_error("psf_measured: no usable image (all peaks too close to an edge)"); return SR459_EINVAL; }
    // run_sr.py:137-150 in float64
    std::vector<double> k((size_t)K * K);
    for (int u = 0; u < K; ++u)
        for (int v = 0; v < K; ++v)
            k[u * K + v] = patch[(size_t)(margin - halfwidth + u) * side + (margin - halfwidth + v)] / (double)used;
    double corners[36]; int c = 0;
    for (int u = 0; u < 3; ++u) for (int v = 0; v < 3; ++v) corners[c++] = k[u * K + v];
    for (int u = 0; u < 3; ++u) for (int v = K - 3; v < K; ++v) corners[c++] = k[u * K + v];
    for (int u = K - 3; u < K; ++u) for (int v = 0; v < 3; ++v) corners[c++] = k[u * K + v];
    for (int u = K - 3; u < K; ++u) for (int v = K - 3; v < K; ++v) corners[c++] = k[u * K + v];
    double bg = np_pairwise_sum(corners, 36) / 36.0;
    for (auto& x : k) { x -= bg; if (x < 0.0) x = 0.0; }
    double s = np_pairwise_sum(k.data(), k.size());
    for (size_t i = 0; i < k.size(); ++i) h_out[i] = k[i] / s;
    return SR459_OK;
}

int sr459_u8_to_f32(const uint8_t* d_in, int64_t n, float* d_out, void* stream) {
    if (!d_in || !d_out || n < 0) { set_error("u8_to_f32: bad arguments"); return SR459_EINVAL; }
    return launch_u8_to_f32(d_in, n, d_out, (cudaStream_t)stream);
}
int sr459_f32_to_u8(const float* d_in, int64_t n, uint8_t* d_out, void* stream) {
    if (!d_in || !d_out || n < 0) { set_error("f32_to_u8: bad arguments"); return SR459_EINVAL; }
    return launch_f32_to_u8(d_in, n, d_out, (cudaStream_t)stream);
}
int sr459_extract_red(const void* d_in, int is_f32, int H, int W, int row_off, int col_off, float* d_out, void* stream) {
    if (!d_in || !d_out || H < 1 || W < 1 || row_off < 0 || row_off > 1 || col_off < 0 || col_off > 1) { set_error("extract_red: bad arguments"); return SR459_EINVAL; }
    return launch_extract_red(d_in, is_f32, H, W, row_off, col_off, d_out, (cudaStream_t)stream);
}
int sr459_mean_frames(const float* const* h_d_frames, int n_frames, int64_t n_px, float* d_out, void* stream) {
    if (!h_d_frames || !d_out || n_px < 0) { set_error("mean_frames: bad arguments"); return SR459_EINVAL; }
    return launch_mean_frames(h_d_frames, n_frames, n_px, d_out, (cudaStream_t)stream);
}

int sr459_ingest_groups_v(const uint8_t* const* h_d_imgs, int n_groups, const int* h_reps, int H, int W, int step, int row_off, int col_off,
                          float* const* h_d_frames, float* d_mean_f32, uint8_t* d_mean_u8, void* stream) {
    if (!h_d_imgs || !h_d_frames || !h_reps) { set_error("ingest_groups: null pointer arrays"); return SR459_EINVAL; }
    if (n_groups < 1 || n_groups > kMaxFrames) { set_error("ingest_groups: 1..%d frames, got %d", kMaxFrames, n_groups); return SR459_EINVAL; }
    int total = 0;
    for (int g = 0; g < n_groups; ++g) {
        if (h_reps[g] < 1) { set_error("ingest_groups: frame %d has %d repetitions", g, h_reps[g]); return SR459_EINVAL; }
        if (!h_d_frames[g]) { set_error("ingest_groups: frame %d is null", g); return SR459_EINVAL; }
        total += h_reps[g];
    }
    for (int k = 0; k < total; ++k) if (!h_d_imgs[k]) { set_error("ingest_groups: image %d is null", k); return SR459_EINVAL; }
    return launch_ingest_groups(h_d_imgs, n_groups, h_reps, H, W, step, row_off, col_off, h_d_frames, d_mean_f32, d_mean_u8, (cudaStream_t)stream);
}

int sr459_ingest_groups(const uint8_t* const* h_d_imgs, int n_groups, int reps, int H, int W, int step, int row_off, int col_off,
                        float* const* h_d_frames, float* d_mean_f32, uint8_t* d_mean_u8, void* stream) {
    int r[kMaxFrames];
    for (int g = 0; g < kMaxFrames; ++g) r[g] = reps;
    return sr459_ingest_groups_v(h_d_imgs, n_groups, r, H, W, step, row_off, col_off, h_d_frames, d_mean_f32, d_mean_u8, stream);
}

int sr459_blur(const float* d_img, int H, int W, const double* h_psf, int kh, int kw, float* d_out, void* stream) {
    if (!d_img || !d_out || !h_psf || H < 1 || W < 1 || kh < 1 || kw < 1) { set_error("blur: bad arguments"); return SR459_EINVAL; }
    if (d_img == d_out) { set_error("blur: in-place operation is not supported"); return SR459_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    TapList t;
    build_taps_host(h_psf, kh, kw, false, t);
    SR_TRY(upload_taps(t));
    CorrArgs c{d_img, H, W, &t, d_out, 0, nullptr, 0.f, 0.f, 0.f};
    int rc = launch_corr2d(c, st);
    cudaStreamSynchronize(st);           // the tap list is freed below
    free_taps(t);
    return rc;
}

int sr459_zoom_to(const float* d_in, int h, int w, int H, int W, int order, float* d_out, void* stream) {
    if (!d_in || !d_out || h < 1 || w < 1 || H < 1 || W < 1) { set_error("zoom: bad arguments"); return SR459_EINVAL; }
    if (order != 0 && order != 1 && order != 3) { set_error("zoom: spline order %d is not implemented (0, 1 and 3 are)", order); return SR459_EUNSUPPORTED; }
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<AxisTap> ty, tx;
    build_zoom_axis(h, H, ty, order);
    build_zoom_axis(w, W, tx, order);
    AxisTap *dty = nullptr, *dtx = nullptr; float* coef = nullptr;
    int rc = upload_taps_tbl(ty, &dty);
    if (rc == SR459_OK) rc = upload_taps_tbl(tx, &dtx);
    if (rc == SR459_OK && order == 3) rc = dev_alloc(&coef, (size_t)h * w);
    if (rc == SR459_OK && order == 3) {
        PrefilterJob j{d_in, coef, h, w, h, w, h, w, SRC_PLAIN, 1};
        rc = launch_prefilter2d(&j, 1, st);
    }
    if (rc == SR459_OK) {
        ResampleSumArgs a{};
        a.coef[0] = order == 3 ? coef : d_in; a.cpitch[0] = w; a.n = 1; a.ty = dty; a.tx = dtx; a.Hout = H; a.Wout = W; a.scale = 1.0f; a.out = d_out;
        rc = launch_resample_sum(a, st);
    }
    cudaStreamSynchronize(st);
    if (dty) gfree(dty); if (dtx) gfree(dtx); if (coef) gfree(coef);
    return rc;
}

int sr459_zoom(const float* d_in, int h, int w, int factor, float* d_out, void* stream) {
    if (factor < 1) { set_error("zoom: bad arguments"); return SR459_EINVAL; }
    return sr459_zoom_to(d_in, h, w, h * factor, w * factor, 3, d_out, stream);
}

// shared body of shift / forward_model / back_project: out = spline(pad(src virtual)) evaluated with tables
static int shift_like(const float* d_src, int sh, int sw, int vh, int vw, int mode, int factor, double sy, double sx,
                      int oh, int ow, int stride, float* d_out, cudaStream_t st) {
    const int PH = vh + 2 * kPad, PW = vw + 2 * kPad;
    std::vector<AxisTap> ty, tx;
    build_shift_axis(vh, sy, oh, stride, ty);
    build_shift_axis(vw, sx, ow, stride, tx);
    AxisTap *dty = nullptr, *dtx = nullptr; float* coef = nullptr;
    int rc = upload_taps_tbl(ty, &dty);
    if (rc == SR459_OK) rc = upload_taps_tbl(tx, &dtx);
    if (rc == SR459_OK) rc = dev_alloc(&coef, (size_t)PH * PW);
    if (rc == SR459_OK) {
        PrefilterJob j{d_src, coef, sh, sw, vh, vw, PH, PW, mode, factor};
        rc = launch_prefilter2d(&j, 1, st);
    }
    if (rc == SR459_OK) {
        ResampleSumArgs a{};
        a.coef[0] = coef; a.cpitch[0] = PW; a.n = 1; a.ty = dty; a.tx = dtx; a.Hout = oh; a.Wout = ow; a.scale = 1.0f; a.out = d_out;
        rc = launch_resample_sum(a, st);
    }
    cudaStreamSynchronize(st);
    if (dty) gfree(dty); if (dtx) gfree(dtx); if (coef) gfree(coef);
    return rc;
}

int sr459_shift(const float* d_in, int H, int W, double sy, double sx, float* d_out, void* stream) {
    if (!d_in || !d_out || H < 1 || W < 1) { set_error("shift: bad arguments"); return SR459_EINVAL; }
    return shift_like(d_in, H, W, H, W, SRC_EDGEPAD, 1, sy, sx, H, W, 1, d_out, (cudaStream_t)stream);
}

int sr459_forward_model(const float* d_hr, int H, int W, const double* h_psf, int kh, int kw, double dy, double dx,
                        int factor, float* d_out_lr, void* stream) {
    if (!d_hr || !d_out_lr || !h_psf || H < 1 || W < 1 || factor < 1) { set_error("forward_model: bad arguments"); return SR459_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    float* B = nullptr;
    SR_TRY(dev_alloc(&B, (size_t)H * W));
    int rc = sr459_blur(d_hr, H, W, h_psf, kh, kw, B, stream);
    const int oh = (H + factor - 1) / factor, ow = (W + factor - 1) / factor;
    if (rc == SR459_OK) rc = shift_like(B, H, W, H, W, SRC_EDGEPAD, 1, dy * factor, dx * factor, oh, ow, factor, d_out_lr, st);
    gfree(B);
    return rc;
}

int sr459_back_project(const float* d_err, int h, int w, const double* h_psf, int kh, int kw, double dy, double dx,
                       int factor, int H, int W, float* d_out_hr, void* stream) {
    if (!d_err || !d_out_hr || !h_psf || h < 1 || w < 1 || H < 1 || W < 1 || factor < 1) { set_error("back_project: bad arguments"); return SR459_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    float* S = nullptr;
    SR_TRY(dev_alloc(&S, (size_t)H * W));
    int rc = shift_like(d_err, h, w, H, W, SRC_ZI_EDGEPAD, factor, -dy * factor, -dx * factor, H, W, 1, S, st);
    if (rc == SR459_OK) {
        TapList t;
        build_taps_host(h_psf, kh, kw, true, t);
        rc = upload_taps(t);
        if (rc == SR459_OK) {
            CorrArgs c{S, H, W, &t, d_out_hr, 0, nullptr, 0.f, 0.f, 0.f};
            rc = launch_corr2d(c, st);
        }
        cudaStreamSynchronize(st);
        free_taps(t);
    }
    gfree(S);
    return rc;
}

int sr459_plan_create(sr459_plan** out, int device, int n_frames, int h, int w, int factor, const double* h_shifts_yx,
                      const double* h_psf, int kh, int kw, double step, double clip_lo, double clip_hi) {
    if (!out) { set_error("plan_create: null output"); return SR459_EINVAL; }
    *out = nullptr;
    SR_TRY(plan_args_check(n_frames, h, w, factor, h_shifts_yx, h_psf, kh, kw));
    SR_DEVICE(device);
    sr459_plan* p = new (std::nothrow) sr459_plan();
    if (!p) { set_error("out of host memory"); return SR459_ENOMEM; }
    int rc = plan_common_init(p, device, n_frames, h, w, factor, h_shifts_yx, h_psf, kh, kw, step, clip_lo, clip_hi);
    if (rc != SR459_OK) { plan_free(p); return rc; }
    *out = p;
    return SR459_OK;
}

void sr459_plan_destroy(sr459_plan* plan) { plan_free(plan); }

int sr459_plan_is_integer(const sr459_plan* plan) { return plan && plan->is_int ? 1 : 0; }

int sr459_plan_set_path(sr459_plan* plan, int path) {
    if (!plan || path < 0 || path > 8) { set_error("plan_set_path: bad arguments"); return SR459_EINVAL; }
    if (path == 8 && !plan->lat) { set_error("lattice kernel not available for this plan (needs integer |shift| <= 1, factor 2, PSF reach <= 15, W % 4 == 0)"); return SR459_EUNSUPPORTED; }
    if (path == 5 || path == 6) { set_error("paths 5 and 6 (the four-phase and the persistent separable kernels) were removed; use 4 (k_ibp_ws) or 7 (streaming)"); return SR459_EUNSUPPORTED; }
    if (path >= 4 && path <= 7 && !(fused_supported(plan) && ibp_sep_supported(plan->kh, plan->kw, plan->smax, plan->separable, plan->W))) {
        set_error("separable TMA kernel not available for this plan (needs integer |shift| <= 1, a rank-1 PSF up to 7x7, W % 4 == 0)");
        return SR459_EUNSUPPORTED;
    }
    if (path == 7 && (plan->band || !(ibp_stream_supported(plan->kh, plan->kw, plan->smax, plan->separable, plan->W, plan->nslot) ||
                                      ibp_stream_ml_supported(plan->kh, plan->kw, plan->smax, plan->separable, plan->W, plan->nslot)))) {
        set_error("streaming kernel not available for this plan (needs a whole-image plan)");
        return SR459_EUNSUPPORTED;
    }
    if ((path == 2 || path == 3) && !fused_supported(plan)) { set_error("fused integer path not available for this plan"); return SR459_EUNSUPPORTED; }
    if (path == 3 && !(ibp_fast_supported(plan->kh, plan->kw, plan->smax) && plan->W % 4 == 0)) { set_error("fast interior kernel not available for this plan"); return SR459_EUNSUPPORTED; }
    plan->path = path;
    return SR459_OK;
}

int sr459_fir_bp_axis(int n_lr, int factor, double shift, int* h_i0, float* h_w) {
    if (n_lr < 1 || factor < 1 || !h_i0 || !h_w) { set_error("fir_bp_axis: bad arguments"); return SR459_EINVAL; }
    std::vector<int> i0; std::vector<float> w;
    if (!build_fir_bp_axis(n_lr, factor, shift, i0, w)) { set_error("fir_bp_axis: axis of %d samples is shorter than %d taps (or the shift is too large)", n_lr, kFirTaps); return SR459_EUNSUPPORTED; }
    std::copy(i0.begin(), i0.end(), h_i0);
    std::copy(w.begin(), w.end(), h_w);
    return SR459_OK;
}

int64_t sr459_plan_scratch_bytes(const sr459_plan* plan) { return plan ? plan->scratch + lat_scratch_bytes(plan->lat) : 0; }

int sr459_native2x(sr459_plan* plan, const float* d_mean_lr, float* d_hr, void* stream) {
    if (!plan || !d_mean_lr || !d_hr) { set_error("native2x: bad arguments"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    return run_zoom(plan, d_mean_lr, d_hr, (cudaStream_t)stream);
}

int sr459_saa(sr459_plan* plan, const float* const* h_d_lr, float* d_hr, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr));
    if (!d_hr) { set_error("saa: null output"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    return run_saa(plan, h_d_lr, d_hr, (cudaStream_t)stream);
}

int sr459_ibp_errors(sr459_plan* plan, int n_iter, double* h_errors, void* stream) {
    if (!plan || !h_errors || n_iter < 0 || n_iter > plan->errors_cap) { set_error("ibp_errors: bad arguments"); return SR459_EINVAL; }
    if (n_iter == 0) return SR459_OK;
    cudaStream_t st = (cudaStream_t)stream;
    SR_CUDA(cudaMemcpyAsync(h_errors, plan->d_errors, sizeof(double) * n_iter, cudaMemcpyDeviceToHost, st));
    SR_CUDA(cudaStreamSynchronize(st));
    return SR459_OK;
}

int sr459_ibp(sr459_plan* plan, const float* const* h_d_lr, float* d_hr_inout, int n_iter, double* h_errors, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr));
    if (!d_hr_inout) { set_error("ibp: null hr buffer"); return SR459_EINVAL; }
    if (plan->band) { set_error("ibp: use sr459_band_iterate with a band plan"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    SR_TRY(run_ibp(plan, h_d_lr, d_hr_inout, n_iter, (cudaStream_t)stream));
    if (h_errors) return sr459_ibp_errors(plan, n_iter, h_errors, stream);
    return SR459_OK;
}

int sr459_saa_ibp(sr459_plan* plan, const float* const* h_d_lr, float* d_hr, float* d_saa, int n_iter, double* h_errors, void* stream) {
    SR_TRY(sr459_saa(plan, h_d_lr, d_hr, stream));
    if (d_saa)
        SR_CUDA(cudaMemcpyAsync(d_saa, d_hr, sizeof(float) * (size_t)plan->H * plan->W, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return sr459_ibp(plan, h_d_lr, d_hr, n_iter, h_errors, stream);
}

}  // extern "C"
