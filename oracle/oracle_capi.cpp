/*
 * oracle_capi.cpp -- C entry points over Oracle B (oracle/ocean_oracle.hpp) for ctypes.
 * TEST INFRASTRUCTURE ONLY (see the header of ocean_oracle.hpp).
 */
#include "ocean_oracle.hpp"
#include "../include/ocean_noise.h"

#ifdef _OPENMP
#include <omp.h>
#endif

using namespace ocean_oracle;

namespace {
struct Handle {
    bool dbl;
    Ocean<float>* f;
    Ocean<double>* d;
    std::vector<float> rowsF[4];
    std::vector<double> rowsD[4];
};

template <class T>
const std::vector<T>* pick(Ocean<T>* o, std::vector<T>* rows, int which) {
    switch (which) {
        case 0: return &o->h0;
        case 1: return &o->wave;
        case 2: case 3: case 4: case 5: return &o->spec[which - 2];
        case 6: return &o->ping;
        case 7: return &o->disp;
        case 8: return &o->deriv;
        case 9: return &o->turb;
        case 10: return &o->bfly;
        case 11: return &o->noise;
        case 12: case 13: case 14: case 15: return &rows[which - 12];
        default: return nullptr;
    }
}
}  // namespace

extern "C" {

// Noise texture [y][x][2] from the shared counter-based generator (include/ocean_noise.h).
void oracle_fill_noise(uint64_t seed, int N, float* out) {
#pragma omp parallel for schedule(static)
    for (int y = 0; y < N; ++y)
        for (int x = 0; x < N; ++x)
            for (int ch = 0; ch < 2; ++ch)
                out[((size_t)y * N + x) * 2 + ch] = ocean_noise_texel(seed, (uint32_t)N, (uint32_t)x, (uint32_t)y, (uint32_t)ch);
}

// params: windSpeed, windDirX, windDirY, gravity, fetch, depth; cascades: C x (wavelength, cutoffLow, cutoffHigh, swell, fade)
void* oracle_create(int N, int C, const float* params, const float* cascades, const float* noise, int use_double) {
    OceanParams p{params[0], params[1], params[2], params[3], params[4], params[5]};
    std::vector<CascadeParams> c(C);
    for (int i = 0; i < C; ++i)
        c[i] = CascadeParams{cascades[i * 5 + 0], cascades[i * 5 + 1], cascades[i * 5 + 2], cascades[i * 5 + 3], cascades[i * 5 + 4]};
    Handle* h = new Handle{use_double != 0, nullptr, nullptr, {}, {}};
    if (use_double) h->d = new Ocean<double>(N, C, p, c.data(), noise);
    else h->f = new Ocean<float>(N, C, p, c.data(), noise);
    return h;
}

void oracle_destroy(void* hv) {
    Handle* h = (Handle*)hv;
    delete h->f; delete h->d; delete h;
}

void oracle_set_threads(void* hv, int n) {
    Handle* h = (Handle*)hv;
    if (n <= 0) {
#ifdef _OPENMP
        n = omp_get_max_threads();
#else
        n = 1;
#endif
    }
    if (h->dbl) h->d->threads = n; else h->f->threads = n;
}

int oracle_get_threads(void* hv) {
    Handle* h = (Handle*)hv;
    return h->dbl ? h->d->threads : h->f->threads;
}

void oracle_init_spectrum(void* hv) {
    Handle* h = (Handle*)hv;
    if (h->dbl) h->d->init_spectrum(); else h->f->init_spectrum();
}

void oracle_step(void* hv, double time) {
    Handle* h = (Handle*)hv;
    if (h->dbl) h->d->step(time); else h->f->step((float)time);
}

void oracle_evolve(void* hv, double time) {
    Handle* h = (Handle*)hv;
    if (h->dbl) h->d->evolve(time); else h->f->evolve((float)time);
}

// evolve + horizontal passes only; results readable as which = 12..15
void oracle_step_rows_only(void* hv, double time) {
    Handle* h = (Handle*)hv;
    if (h->dbl) h->d->step_rows_only(time, h->rowsD); else h->f->step_rows_only((float)time, h->rowsF);
}

void oracle_reset_foam(void* hv) {
    Handle* h = (Handle*)hv;
    if (h->dbl) std::fill(h->d->turb.begin(), h->d->turb.end(), 0.0);
    else std::fill(h->f->turb.begin(), h->f->turb.end(), 0.0f);
}

void oracle_set_turbulence(void* hv, const float* src) {
    Handle* h = (Handle*)hv;
    if (h->dbl) for (size_t i = 0; i < h->d->turb.size(); ++i) h->d->turb[i] = src[i];
    else for (size_t i = 0; i < h->f->turb.size(); ++i) h->f->turb[i] = src[i];
}

long long oracle_size(void* hv, int which) {
    Handle* h = (Handle*)hv;
    if (h->dbl) { auto* v = pick(h->d, h->rowsD, which); return v ? (long long)v->size() : -1; }
    auto* v = pick(h->f, h->rowsF, which); return v ? (long long)v->size() : -1;
}

int oracle_get_f64(void* hv, int which, double* out) {
    Handle* h = (Handle*)hv;
    if (h->dbl) {
        auto* v = pick(h->d, h->rowsD, which); if (!v) return -1;
        std::memcpy(out, v->data(), v->size() * sizeof(double));
    } else {
        auto* v = pick(h->f, h->rowsF, which); if (!v) return -1;
        for (size_t i = 0; i < v->size(); ++i) out[i] = (*v)[i];
    }
    return 0;
}

int oracle_get_f32(void* hv, int which, float* out) {
    Handle* h = (Handle*)hv;
    if (h->dbl) {
        auto* v = pick(h->d, h->rowsD, which); if (!v) return -1;
        for (size_t i = 0; i < v->size(); ++i) out[i] = (float)(*v)[i];
    } else {
        auto* v = pick(h->f, h->rowsF, which); if (!v) return -1;
        std::memcpy(out, v->data(), v->size() * sizeof(float));
    }
    return 0;
}

void oracle_query_heights(void* hv, const float* xz, int n, float* out) {
    Handle* h = (Handle*)hv;
    for (int i = 0; i < n; ++i)
        out[i] = h->dbl ? (float)h->d->water_height(xz[2 * i], xz[2 * i + 1])
                        : h->f->water_height(xz[2 * i], xz[2 * i + 1]);
}

}  // extern "C"
