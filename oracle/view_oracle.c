/*
 * view_oracle.c -- CPU restatement of MV3D's LiDAR view encoder (TEST INFRASTRUCTURE ONLY).
 *
 * This file is the parity oracle for the CUDA path in mv3d_b200/csrc. It is NOT part of the
 * product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it. The product path never falls back to it.
 *
 * It restates, with flat arrays and no replay logs, what the reference routines compute:
 *   createTopMaps          /root/reference/src/lidar_data_preprocess/Python_to_C_Interface/ver3/LidarTopPreprocess.c:30-243
 *   createTopAndFrontMaps  .../ver3/LidarTopAndFrontPreprocess.c:30-295
 *
 * Parity status: PINNED by execution. The reference ships no golden vectors for this path
 * (SURVEY.md section 8c); the restatement is pinned against the reference's own unmodified C sources
 * compiled into oracle/_ref/ (oracle/Makefile) -- tests/test_oracle_vs_reference.py compares
 * them bit-for-bit on seeded sweeps and the Q1-Q12 quirk cases, and tests/golden/ holds outputs
 * of that reference build for boxes where /root/reference is absent.
 *
 * Arithmetic notes (each mirrors a reference line; float means IEEE binary32, one rounding per op):
 *   - cell index: q = (p - MIN) / DIV in float, two separately rounded ops
 *       LidarTopPreprocess.c:107-109 truncates (int)q; LidarTopAndFrontPreprocess.c:124-126 uses (int)floor(q).
 *   - gate: X>=1 && Y>=1 && Z>=0 && X<X_SIZE && Y<Y_SIZE && Z<Z_SIZE      (:112 / :129)
 *   - slice height: strict-greater max of h = z - z_MIN over a zero-initialised map (:115-129 / :133-147);
 *       emitted as h / z_DIV - k                                            (:205 / :143,258)
 *   - density: one count per gated point (:131-138 / :150-156); emitted as
 *       (float)((double)logf(count + 1.0f) / log(32.0)), clamped to 1       (:167-172 / :220-225)
 *   - intensity: intensity of the first point (input order) reaching the cell's max h (:141-152 / :159-171)
 *   - output cell is (X-1, Y-1); index Y_SIZE*(Z_SIZE+2)*(X-1) + (Z_SIZE+2)*(Y-1) + plane (:200,220,236)
 *   - front view (LidarTopAndFrontPreprocess.c:36-38,173-193), RAD_TO_DEG = (float)(180/3.141596), truncating
 *       (int), inclusive upper bounds, plain stores (last point in input order wins). The two bins have DIFFERENT
 *       types (the file is C++: <math.h> is libstdc++'s wrapper, `using namespace std`):
 *         row R (:173): pow(float, int) promotes to double, so sqrt/atan2 and the scaling are all double;
 *         col C (:174): atan2(point.y, point.x) has two float arguments -> std::atan2(float, float) = atan2f,
 *                       and `* RAD_TO_DEG / theta_DIVISION` stays in float (cvttss2si). `nm -D` on the reference
 *                       build imports both atan2 and atan2f. tests/test_oracle_vs_reference.py compares the
 *                       per-point pixel with the reference build one point per call.
 *       The reference writes past the end of front_data when
 *       r_coord == R_SIZE; this restatement applies the same linear index and DROPS stores at or
 *       beyond C_SIZE*R_SIZE*F_SIZE, so in-bounds cells stay identical.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC view_oracle.c -o _build/libview_oracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* x86 cvttss2si / cvttsd2si semantics, which is what the reference's (int) casts compile to:
 * NaN and out-of-range inputs yield INT_MIN ("integer indefinite"). Written out so the oracle
 * does not depend on undefined behaviour of the C cast. */
static int cvtt_f32(float v)
{
    if (!(v > -2147483904.0f && v < 2147483648.0f)) return INT32_MIN;
    return (int)v;
}

static int cvtt_f64(double v)
{
    if (!(v > -2147483649.0 && v < 2147483648.0)) return INT32_MIN;
    return (int)v;
}

/* Density normalisation with the reference's exact type mix: logf on a float argument,
 * double log(32), quotient in double, narrowed to float, then clamped. */
static float density_value(float count)
{
    volatile float arg = count + 1.0f;
    volatile double base = 32.0;
    float val = (float)((double)logf(arg) / log(base));
    return (val < 1.0f) ? val : 1.0f;
}

enum { QUANT_TRUNC = 0, QUANT_FLOOR = 1 };

typedef struct {
    float x_min, x_max, y_min, y_max, z_min, z_max, x_div, y_div, z_div;
    int xn, yn, zn;
} top_params_t;

typedef struct {
    float phi_div, theta_div;
    float c_size, r_size, f_size;
} front_params_t;

static int quantise(float p, float lo, float div, int mode)
{
    float q = (p - lo) / div;
    if (mode == QUANT_FLOOR) q = floorf(q);
    return cvtt_f32(q);
}

/* Core: one pass over the points into dense scratch, then one pass over the cells.
 * top may be NULL (front view only); front may be NULL (bird's-eye view only). */
static int encode(const float *pts, int num, float *top, float *front,
                  const top_params_t *tp, const front_params_t *fp, int mode)
{
    const int xn = tp->xn, yn = tp->yn, zn = tp->zn;
    const size_t cells = (size_t)(xn > 0 ? xn : 0) * (size_t)(yn > 0 ? yn : 0);
    const size_t planes = (size_t)zn + 2;
    float *slice_h = NULL, *cell_h = NULL, *cell_i = NULL, *cell_n = NULL;

    if (top) {
        slice_h = (float *)calloc(cells * (size_t)(zn > 0 ? zn : 0) + 1, sizeof(float));
        cell_h = (float *)calloc(cells + 1, sizeof(float));
        cell_i = (float *)calloc(cells + 1, sizeof(float));
        cell_n = (float *)calloc(cells + 1, sizeof(float));
        if (!slice_h || !cell_h || !cell_i || !cell_n) {
            free(slice_h); free(cell_h); free(cell_i); free(cell_n);
            return -1;
        }
    }

    float rad_to_deg = 0.0f;
    int center_c = 0, center_r = 0;
    long front_len = 0;
    if (front) {
        rad_to_deg = (float)(180 / 3.141596);
        center_c = cvtt_f32(fp->c_size / 2);
        center_r = cvtt_f32(fp->r_size / 2);
        front_len = cvtt_f32(fp->c_size * fp->r_size * fp->f_size);
        for (long i = 0; i < front_len; i++) front[i] = 0.0f;
    }

    for (int i = 0; i < num; i++) {
        const float x = pts[4 * (size_t)i + 0];
        const float y = pts[4 * (size_t)i + 1];
        const float z = pts[4 * (size_t)i + 2];
        const float w = pts[4 * (size_t)i + 3];

        const int X = quantise(x, tp->x_min, tp->x_div, mode);
        const int Y = quantise(y, tp->y_min, tp->y_div, mode);
        const int Z = quantise(z, tp->z_min, tp->z_div, mode);
        if (!(X >= 1 && Y >= 1 && Z >= 0 && X < xn && Y < yn && Z < zn)) continue;

        if (top) {
            const float h = z - tp->z_min;
            const size_t cell = (size_t)X * yn + Y;
            float *slot = &slice_h[cell * zn + Z];
            if (h > *slot) *slot = h;
            cell_n[cell] += 1.0f; /* float counter, as in the reference (saturates at 2^24) */
            if (h > cell_h[cell]) {
                cell_h[cell] = h;
                cell_i[cell] = w;
            }
        }

        if (front) {
            const double dx = x, dy = y, dz = z;
            const double planar = sqrt(dx * dx + dy * dy);
            const int R = cvtt_f64(atan2(dz, planar) * (double)rad_to_deg / (double)fp->phi_div);
            const int C = cvtt_f32(atan2f(y, x) * rad_to_deg / fp->theta_div); /* float: atan2(float, float) */
            /* int + int compared against a float bound, as in the reference */
            const long r_coord = (long)R + center_r, c_coord = (long)C + center_c;
            if ((float)r_coord <= fp->r_size && (float)c_coord <= fp->c_size && r_coord >= 0 && c_coord >= 0) {
                const float base = fp->c_size * fp->f_size * (float)r_coord + fp->f_size * (float)c_coord;
                const int v0 = cvtt_f32((z - tp->z_min) / (tp->z_max - tp->z_min) * 255.0f);
                const int v1 = cvtt_f64(planar / 120.0 * 255.0);
                const int v2 = cvtt_f32((w - 0.0f) * 255.0f);
                const int i0 = cvtt_f32(base + 0.0f), i1 = cvtt_f32(base + 1.0f), i2 = cvtt_f32(base + 2.0f);
                if (i0 >= 0 && i0 < front_len) front[i0] = (float)v0;
                if (i1 >= 0 && i1 < front_len) front[i1] = (float)v1;
                if (i2 >= 0 && i2 < front_len) front[i2] = (float)v2;
            }
        }
    }

    if (top) {
        memset(top, 0, cells * planes * sizeof(float));
        for (int X = 1; X < xn; X++) {
            for (int Y = 1; Y < yn; Y++) {
                const size_t cell = (size_t)X * yn + Y;
                if (cell_n[cell] == 0.0f) continue;
                float *out = top + ((size_t)(X - 1) * yn + (size_t)(Y - 1)) * planes;
                for (int k = 0; k < zn; k++) {
                    const float h = slice_h[cell * zn + k];
                    if (h > 0.0f) out[k] = h / tp->z_div - (float)k;
                }
                if (cell_h[cell] > 0.0f) out[zn] = cell_i[cell];
                out[zn + 1] = density_value(cell_n[cell]);
            }
        }
        free(slice_h); free(cell_h); free(cell_i); free(cell_n);
    }
    return 0;
}

/* Same argument list as the reference's createTopMaps (LidarTopPreprocess.c:30). */
int oracle_createTopMaps(const void *raw_data, int num, void *top_data,
                         float x_MIN, float x_MAX, float y_MIN, float y_MAX, float z_MIN, float z_MAX,
                         float x_DIVISION, float y_DIVISION, float z_DIVISION,
                         int X_SIZE, int Y_SIZE, int Z_SIZE)
{
    top_params_t tp = { x_MIN, x_MAX, y_MIN, y_MAX, z_MIN, z_MAX, x_DIVISION, y_DIVISION, z_DIVISION,
                        X_SIZE, Y_SIZE, Z_SIZE };
    return encode((const float *)raw_data, num, (float *)top_data, NULL, &tp, NULL, QUANT_TRUNC);
}

/* Same argument list as the reference's createTopAndFrontMaps (LidarTopAndFrontPreprocess.c:30).
 * top_data or front_data may be NULL to skip that view (the reference has no such mode; the
 * values of the view that is produced are unaffected). */
int oracle_createTopAndFrontMaps(const void *raw_data, int num, void *top_data, void *front_data,
                                 float x_MIN, float x_MAX, float y_MIN, float y_MAX, float z_MIN, float z_MAX,
                                 float x_DIVISION, float y_DIVISION, float z_DIVISION,
                                 int X_SIZE, int Y_SIZE, int Z_SIZE,
                                 float phi_MIN, float phi_MAX, float theta_MIN, float theta_MAX,
                                 float phi_DIVISION, float theta_DIVISION,
                                 float C_SIZE, float R_SIZE, float F_SIZE)
{
    (void)phi_MIN; (void)phi_MAX; (void)theta_MIN; (void)theta_MAX; /* dead in the reference too */
    top_params_t tp = { x_MIN, x_MAX, y_MIN, y_MAX, z_MIN, z_MAX, x_DIVISION, y_DIVISION, z_DIVISION,
                        X_SIZE, Y_SIZE, Z_SIZE };
    front_params_t fp = { phi_DIVISION, theta_DIVISION, C_SIZE, R_SIZE, F_SIZE };
    return encode((const float *)raw_data, num, (float *)top_data, (float *)front_data, &tp, &fp, QUANT_FLOOR);
}

/* Per-point decisions (test hook): cell[i] = (X-1)*yn + (Y-1) or -1 when the gate rejects point i;
 * pix[i] = r_coord*C + c_coord of the front-view store, or -1 when the reference would not store in bounds.
 * mode: 0 = createTopMaps quantisation (truncate), 1 = createTopAndFrontMaps (floor). pix may be NULL. */
void oracle_point_keys(const void *raw_data, int num, int mode,
                       float x_MIN, float y_MIN, float z_MIN, float x_DIVISION, float y_DIVISION, float z_DIVISION,
                       int X_SIZE, int Y_SIZE, int Z_SIZE, float phi_DIVISION, float theta_DIVISION,
                       float C_SIZE, float R_SIZE, int *cell, int *pix)
{
    const float *pts = (const float *)raw_data;
    const float rad_to_deg = (float)(180 / 3.141596);
    const int center_c = cvtt_f32(C_SIZE / 2), center_r = cvtt_f32(R_SIZE / 2);
    const long npix = (long)cvtt_f32(C_SIZE) * (long)cvtt_f32(R_SIZE);
    for (int i = 0; i < num; i++) {
        const float x = pts[4 * (size_t)i], y = pts[4 * (size_t)i + 1], z = pts[4 * (size_t)i + 2];
        const int X = quantise(x, x_MIN, x_DIVISION, mode);
        const int Y = quantise(y, y_MIN, y_DIVISION, mode);
        const int Z = quantise(z, z_MIN, z_DIVISION, mode);
        const int keep = (X >= 1 && Y >= 1 && Z >= 0 && X < X_SIZE && Y < Y_SIZE && Z < Z_SIZE);
        cell[i] = keep ? (X - 1) * Y_SIZE + (Y - 1) : -1;
        if (!pix) continue;
        pix[i] = -1;
        if (!keep) continue;
        const double dx = x, dy = y, dz = z;
        const int R = cvtt_f64(atan2(dz, sqrt(dx * dx + dy * dy)) * (double)rad_to_deg / (double)phi_DIVISION);
        const int C = cvtt_f32(atan2f(y, x) * rad_to_deg / theta_DIVISION);
        const long r = (long)R + center_r, c = (long)C + center_c;
        if ((float)r <= R_SIZE && (float)c <= C_SIZE && r >= 0 && c >= 0) {
            const long q = r * (long)cvtt_f32(C_SIZE) + c;
            if (q < npix) pix[i] = (int)q;
        }
    }
}

/* The 32 distinct density values (count 0..31; anything above is 1.0f), for host-side tests. */
void oracle_density_table(float *out32)
{
    for (int c = 0; c < 32; c++) out32[c] = density_value((float)c);
}
