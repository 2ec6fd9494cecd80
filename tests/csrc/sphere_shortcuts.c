/* CPU sweep for the exact shortcuts of sphere_t (raytracingrust_b200/csrc/device_core.cuh): the reference sequence
 * (Sphere::hit, lib/objects.rs:23-44) against the shortcut form, bit for bit, on random and adversarial inputs.
 * Build: gcc -O2 -ffp-contract=off (no FMA contraction, as the kernels' -fmad=false). Prints mismatches. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

typedef struct { float x, y, z; } v3;
static float dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }

static int plain(v3 o, v3 d, v3 ctr, float rad, float tmin, float tmax, float* t) {
    v3 oc = {o.x - ctr.x, o.y - ctr.y, o.z - ctr.z};
    float a = dot(d, d), half_b = dot(oc, d), c = dot(oc, oc) - rad * rad;
    float disc = half_b * half_b - a * c;
    if (disc < 0.0f) return 0;
    float sqrtd = sqrtf(disc);
    float root = (-half_b - sqrtd) / a;
    if (root < tmin || tmax < root) {
        root = (-half_b + sqrtd) / a;
        if (root < tmin || tmax < root) return 0;
    }
    *t = root;
    return 1;
}
static float div_or_zero(float num, float a) { return (num == 0.0f && a > 0.0f) ? num : num / a; }
static int quick(v3 o, v3 d, v3 ctr, float rad, float tmin, float tmax, float* t) {
    v3 oc = {o.x - ctr.x, o.y - ctr.y, o.z - ctr.z};
    float a = dot(d, d), half_b = dot(oc, d), c = dot(oc, oc) - rad * rad;
    float hb2 = half_b * half_b;
    float disc = hb2 - a * c;
    if (disc < 0.0f) return 0;
    const int tame = disc >= 0.0f && a > 0.0f && tmin > 0.0f && hb2 >= 1.17549435e-38f && hb2 < INFINITY;
    if (tame && c > 0.0f && half_b > 0.0f) return 0;
    float sqrtd = sqrtf(disc);
    int second = tame && c < 0.0f;
    float root = 0.0f;
    if (!second) {
        root = div_or_zero(-half_b - sqrtd, a);
        second = root < tmin || tmax < root;
    }
    if (second) {
        root = div_or_zero(-half_b + sqrtd, a);
        if (root < tmin || tmax < root) return 0;
    }
    *t = root;
    return 1;
}

static uint64_t s_rng = 0x9E3779B97F4A7C15ull;
static uint32_t rnd(void) { s_rng ^= s_rng << 13; s_rng ^= s_rng >> 7; s_rng ^= s_rng << 17; return (uint32_t)(s_rng >> 16); }
static float uni(float lo, float hi) { return lo + (hi - lo) * (float)(rnd() >> 8) * (1.0f / 16777216.0f); }
static float weird(void) {
    static const float W[] = {0.0f, -0.0f, 1.0f, -1.0f, 1e-38f, 1e-45f, 1e-20f, -1e-20f, 1e19f, -1e19f, 3e38f, INFINITY,
                              -INFINITY, NAN, 1e-5f, 100.0f, 0.5f, 1e-10f, -1e-10f, 2.0f};
    return W[rnd() % (sizeof(W) / sizeof(W[0]))];
}
static float bitsf(void) { uint32_t b = rnd() ^ (rnd() << 16); float f; memcpy(&f, &b, 4); return f; }

int main(int argc, char** argv) {
    long n = argc > 1 ? atol(argv[1]) : 20000000, bad = 0, hits = 0, fast = 0;
    (void)fast;
    for (long i = 0; i < n; ++i) {
        v3 o, d, ctr; float rad, tmin = 1e-5f, tmax = INFINITY;
        const int mode = (int)(i % 6);
        ctr.x = uni(-5, 5); ctr.y = uni(-5, 5); ctr.z = uni(-5, 5); rad = uni(0.01f, 3.0f);
        if (i % 11 == 0) { rad = 100.0f; ctr.y -= 100.0f; }
        d.x = uni(-1, 1); d.y = uni(-1, 1); d.z = uni(-1, 1);
        if (mode == 0) { o.x = uni(-6, 6); o.y = uni(-6, 6); o.z = uni(-6, 6); }
        else if (mode <= 3) {  /* origin ON the sphere (scattered rays): the c == 0 / zero-numerator cases */
            v3 u = {uni(-1, 1), uni(-1, 1), uni(-1, 1)};
            float l = sqrtf(dot(u, u)); if (l == 0.0f) l = 1.0f;
            o.x = ctr.x + u.x / l * rad; o.y = ctr.y + u.y / l * rad; o.z = ctr.z + u.z / l * rad;
            if (mode == 3) { d = u; if (i & 8) { d.x = -d.x; d.y = -d.y; d.z = -d.z; } }
        } else if (mode == 4) { o.x = weird(); o.y = weird(); o.z = weird(); d.x = weird(); d.y = weird(); d.z = weird(); if (i & 16) rad = weird(); }
        else { o.x = bitsf(); o.y = bitsf(); o.z = bitsf(); d.x = bitsf(); d.y = bitsf(); d.z = bitsf(); ctr.x = bitsf(); rad = bitsf(); }
        if (i % 7 == 0) tmax = uni(0.0f, 10.0f);
        if (i % 13 == 0) tmin = -INFINITY;   /* Volume::hit's first boundary query (volume.rs:36) */
        if (i % 17 == 0) tmin = uni(-2.0f, 2.0f);
        float t0 = 0.0f, t1 = 0.0f;
        const int h0 = plain(o, d, ctr, rad, tmin, tmax, &t0), h1 = quick(o, d, ctr, rad, tmin, tmax, &t1);
        uint32_t b0, b1; memcpy(&b0, &t0, 4); memcpy(&b1, &t1, 4);
        hits += h0;
        if (h0 != h1 || (h0 && b0 != b1 && !(t0 != t0 && t1 != t1))) {
            if (bad++ < 10) printf("MISMATCH i=%ld mode=%d plain=%d %a quick=%d %a\n", i, mode, h0, t0, h1, t1);
        }
    }
    printf("cases %ld hits %ld mismatches %ld\n", n, hits, bad);
    return bad != 0;
}
