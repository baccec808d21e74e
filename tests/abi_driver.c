/* abi_driver.c -- a plain C caller of liboceanfft.so, compiled with gcc against include/oceanfft.h
 * ("the ABI is exercised through ctypes and a C test driver").  ctypes re-declares every
 * signature by hand; this file cannot drift from the header: if a prototype changes it stops compiling,
 * if a symbol disappears it stops linking.
 *
 *   abi_driver            link + error-path checks that need no device; exit 0
 *   abi_driver --device   create -> init -> step x3 -> copy -> query -> buoyancy -> update/readback -> destroy on cuda:0
 *
 * Prints one line "abi_driver: ..." per stage; any failed expectation exits non-zero.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "oceanfft.h"

#define EXPECT(cond)                                                                              \
    do {                                                                                          \
        if (!(cond)) {                                                                            \
            fprintf(stderr, "abi_driver: FAILED %s (line %d): %s\n", #cond, __LINE__, ocean_last_error()); \
            return 1;                                                                             \
        }                                                                                         \
    } while (0)

/* every entry point of the header, referenced so that the link fails if one is not exported */
static const void* const kSymbols[] = {
    (const void*)ocean_abi_version, (const void*)ocean_last_error, (const void*)ocean_device_count, (const void*)ocean_create,
    (const void*)ocean_destroy, (const void*)ocean_set_params, (const void*)ocean_init_spectrum, (const void*)ocean_step,
    (const void*)ocean_sync, (const void*)ocean_step_partial, (const void*)ocean_update, (const void*)ocean_set_readback_mode,
    (const void*)ocean_readback_mode_of_latest, (const void*)ocean_readback_ptr, (const void*)ocean_get_output,
    (const void*)ocean_copy_output, (const void*)ocean_get_wavelengths, (const void*)ocean_generate_mips,
    (const void*)ocean_set_auto_mips, (const void*)ocean_mip_levels, (const void*)ocean_get_mips, (const void*)ocean_copy_mip,
    (const void*)ocean_sample_surface, (const void*)ocean_buoyancy, (const void*)ocean_set_turbulence,
    (const void*)ocean_reset_foam, (const void*)ocean_frame_index, (const void*)ocean_query_heights,
    (const void*)ocean_get_stream, (const void*)ocean_set_stream, (const void*)ocean_launch_count, (const void*)ocean_describe,
    (const void*)ocean_ensemble_create, (const void*)ocean_ensemble_destroy, (const void*)ocean_ensemble_run,
    (const void*)ocean_ensemble_get_output, (const void*)ocean_ensemble_copy_output, (const void*)ocean_update_buoyancy,
    (const void*)ocean_get_copy_stream, (const void*)ocean_buoyancy_noalloc,
};

static const OceanCascade kCascades[2] = {{1530.0f, 1e-10f, 1e12f, 0.4f, 0.1f}, {201.0f, 1e-5f, 1e6f, 0.1f, 0.1f}};

static int host_checks(void) {
    OceanHandle* h = (OceanHandle*)0x1;
    OceanParams p = {8.0f, 1.0f, -1.0f, 9.81f, 50000.0f, 2560.0f, 64, 1.0f};
    EXPECT(ocean_abi_version() == OCEAN_ABI_VERSION);
    EXPECT(sizeof(OceanParams) == 32 && sizeof(OceanCascade) == 20 && sizeof(OceanBuoyantBox) == 24);
    /* argument validation happens before any CUDA call */
    EXPECT(ocean_create(NULL, kCascades, 2, NULL, 0, 0, &h) == OCEAN_ERR_INVALID_ARGUMENT && h == NULL);
    p.textureSize = 100;
    EXPECT(ocean_create(&p, kCascades, 2, NULL, 0, 0, &h) == OCEAN_ERR_INVALID_SIZE);
    EXPECT(strstr(ocean_last_error(), "power of two") != NULL);
    p.textureSize = 64;
    EXPECT(ocean_create(&p, kCascades, 6, NULL, 0, 0, &h) == OCEAN_ERR_INVALID_CASCADES);
    p.windDirX = p.windDirY = 0.0f;
    EXPECT(ocean_create(&p, kCascades, 2, NULL, 0, 0, &h) == OCEAN_ERR_INVALID_PARAMS);
    EXPECT(ocean_step(NULL, 0.0f) == OCEAN_ERR_INVALID_ARGUMENT);
    EXPECT(ocean_frame_index(NULL) == -1);
    ocean_destroy(NULL);
    printf("abi_driver: %d symbols linked, host-side error paths ok, %d CUDA device(s)\n",
           (int)(sizeof kSymbols / sizeof kSymbols[0]), ocean_device_count());
    return 0;
}

static int device_flow(void) {
    enum { N = 64, C = 2, NQ = 5 };
    OceanParams p = {8.0f, 1.0f, -1.0f, 9.81f, 50000.0f, 2560.0f, N, 1.0f};
    OceanHandle* h = NULL;
    float* disp = (float*)malloc(sizeof(float) * C * N * N * 4);
    const float pts[NQ * 2] = {0.0f, 0.0f, -32.0f, -32.0f, 31.5f, 7.25f, 1000.0f, -1000.0f, 3.0f, -7.0f};
    float heights[NQ], out4[4];
    OceanBuoyantBox box = {3.0f, -0.25f, -7.0f, 1.0f, 2.0f, 1.0f};
    uint64_t before;
    int64_t frame = -1;
    const float* rb;
    int i, x, y;
    char text[512];
    EXPECT(disp != NULL);
    EXPECT(ocean_device_count() >= 1);
    EXPECT(ocean_create(&p, kCascades, C, NULL, 7, 0, &h) == OCEAN_OK && h != NULL);
    EXPECT(ocean_step(h, 0.0f) == OCEAN_ERR_NOT_INITIALIZED);
    EXPECT(ocean_init_spectrum(h) == OCEAN_OK);
    before = ocean_launch_count();
    for (i = 0; i < 3; ++i) EXPECT(ocean_step(h, (float)i / 60.0f) == OCEAN_OK);
    EXPECT(ocean_sync(h) == OCEAN_OK);
    EXPECT(ocean_launch_count() - before == 6); /* two kernels per frame */
    EXPECT(ocean_frame_index(h) == 3);
    EXPECT(ocean_copy_output(h, OCEAN_DISPLACEMENT, -1, disp) == OCEAN_OK);
    EXPECT(ocean_query_heights(h, pts, NQ, OCEAN_QUERY_NEAREST_REF, heights, 0, 0) == OCEAN_OK);
    for (i = 0; i < NQ; ++i) { /* WB:199-208 restated in C */
        float u = (pts[2 * i] - (float)(-(N / 2))) / ((float)(N / 2) - (float)(-(N / 2)));
        float v = (pts[2 * i + 1] - (float)(-(N / 2))) / ((float)(N / 2) - (float)(-(N / 2)));
        u = u < 0.0f ? 0.0f : (u > 1.0f ? 1.0f : u);
        v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
        x = (int)(u * (float)N);
        y = (int)(v * (float)N);
        x = x > N - 1 ? N - 1 : x;
        y = y > N - 1 ? N - 1 : y;
        EXPECT(heights[i] == disp[((size_t)y * N + x) * 4 + 1]);
    }
    EXPECT(ocean_buoyancy(h, &box, 1, OCEAN_QUERY_NEAREST_REF, 1.0f, -9.81f, out4, 0, 0) == OCEAN_OK);
    EXPECT(out4[0] == heights[4]); /* BO:18: the same lookup at the box position */
    EXPECT(out4[1] == fmaxf(0.0f, out4[0] - box.py));
    EXPECT(ocean_update(h, 0.1f) == OCEAN_OK);
    rb = ocean_readback_ptr(h, 1, &frame);
    EXPECT(rb != NULL && frame == 3);
    EXPECT(ocean_copy_output(h, OCEAN_DISPLACEMENT, 0, disp) == OCEAN_OK);
    EXPECT(memcmp(rb, disp, sizeof(float) * N * N * 4) == 0);
    /* the same answer from the owning GPU, one frame later at the same time (same displacement) */
    EXPECT(ocean_update_buoyancy(h, 0.1f, &box, 1, 1.0f, -9.81f, out4) == OCEAN_OK);
    EXPECT(ocean_sync(h) == OCEAN_OK);
    {
        float again[4];
        EXPECT(ocean_buoyancy(h, &box, 1, OCEAN_QUERY_NEAREST_REF, 1.0f, -9.81f, again, 0, 0) == OCEAN_OK);
        EXPECT(memcmp(again, out4, sizeof again) == 0);
    }
    EXPECT(ocean_describe(h, text, sizeof text) == OCEAN_OK && strstr(text, "N=64") != NULL);
    ocean_destroy(h);
    free(disp);
    printf("abi_driver: device flow ok (%s)\n", text);
    return 0;
}

int main(int argc, char** argv) {
    int rc = host_checks();
    if (rc == 0 && argc > 1 && strcmp(argv[1], "--device") == 0) rc = device_flow();
    return rc;
}
