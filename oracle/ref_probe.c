/*
 * ref_probe.c -- per-point probe of the REFERENCE build (TEST INFRASTRUCTURE ONLY).
 *
 * The front view is a "last point wins" image, so comparing whole images of dense clouds hides single-point
 * differences in the pixel index. This helper calls the reference's own createTopAndFrontMaps
 * (oracle/_ref/LidarTopAndFrontPreprocess.so, built unmodified from
 * /root/reference/src/lidar_data_preprocess/Python_to_C_Interface/ver3/LidarTopAndFrontPreprocess.c:30) with ONE point
 * per call and reports where that point landed: pix[i] = r_coord * C + c_coord of the pixel the reference wrote
 * (LidarTopAndFrontPreprocess.c:173-193), -1 if it wrote nothing inside the R*C*F image, and cell[i] = the
 * bird's-eye-view cell it touched (-1 if gated out). Every point is given intensity 0.5 for the call, so the pixel's
 * third channel is (int)(0.5*255) = 127 and can be found without knowing the index arithmetic.
 *
 * `guess` (optional) is the caller's prediction of pix[i]: the probe looks there first and scans the image only when
 * the prediction is wrong, which keeps a call at the cost of the reference routine itself.
 *
 * Build: gcc -O2 -shared -fPIC ref_probe.c -o _build/libref_probe.so -ldl
 */
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

typedef void (*topfront_fn)(const void *, int, const void *, const void *, float, float, float, float, float, float,
                            float, float, float, int, int, int, float, float, float, float, float, float, float,
                            float, float);

/* returns 0, or -1 when the library / symbol cannot be loaded */
int ref_probe_points(const char *so_path, const float *pts, int n, float x_min, float x_max, float y_min, float y_max,
                     float z_min, float z_max, float x_div, float y_div, float z_div, int xn, int yn, int zn,
                     float phi_div, float theta_div, int cn, int rn, const int *guess, int *cell, int *pix)
{
    void *h = dlopen(so_path, RTLD_NOW | RTLD_LOCAL);
    if (!h) return -1;
    topfront_fn fn = (topfront_fn)dlsym(h, "createTopAndFrontMaps");
    if (!fn) { dlclose(h); return -1; }
    const size_t planes = (size_t)zn + 2, cells = (size_t)xn * yn;
    const long npix = (long)rn * cn;
    /* one row + one pixel of slack behind the image: the reference stores past the end when r_coord == R (Q9) */
    const size_t front_len = (size_t)npix * 3 + (size_t)cn * 3 + 64;
    float *top = (float *)malloc((cells * planes + 1) * sizeof(float));
    float *front = (float *)malloc(front_len * sizeof(float));
    if (!top || !front) { free(top); free(front); dlclose(h); return -1; }
    for (int i = 0; i < n; i++) {
        float p[4] = {pts[4 * (size_t)i], pts[4 * (size_t)i + 1], pts[4 * (size_t)i + 2], 0.5f};
        memset(front + (size_t)npix * 3, 0, (front_len - (size_t)npix * 3) * sizeof(float));
        fn(p, 1, top, front, x_min, x_max, y_min, y_max, z_min, z_max, x_div, y_div, z_div, xn, yn, zn, 0.0f, 0.0f,
           0.0f, 0.0f, phi_div, theta_div, (float)cn, (float)rn, 3.0f);
        cell[i] = -1;
        for (size_t c = 0; c < cells; c++)
            if (top[c * planes + planes - 1] != 0.0f) { cell[i] = (int)c; break; }  /* density > 0: the touched cell */
        pix[i] = -1;
        const long g = guess ? guess[i] : -1;
        if (g >= 0 && g < npix && front[3 * g + 2] == 127.0f) { pix[i] = (int)g; continue; }
        for (long q = 0; q < npix; q++)
            if (front[3 * q + 2] == 127.0f) { pix[i] = (int)q; break; }
    }
    free(top); free(front);
    dlclose(h);
    return 0;
}
