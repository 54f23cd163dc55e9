/* drivers/minifb_headless.c -- implementation of include/shim/MiniFB.h (see there). */
#define _POSIX_C_SOURCE 200809L
#include <MiniFB.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../example-gui-opencl-raytracer_b200/csrc/png_io.h"

#define MAX_KEYS 16
struct mfb_window {
    unsigned w, h;
    mfb_keyboard_func on_key;
    int frames_wanted, frames_done;
    mfb_key keys[MAX_KEYS];
    int nkeys;
    double* stamp; /* time of each mfb_update_ex */
    double t_first_wait;
    const void* last;
};
struct mfb_timer { double t0; };
static struct mfb_window* g_win; /* rayinteractive.c owns one window */

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}
static mfb_key key_by_name(const char* s) {
    static const struct { const char* n; mfb_key k; } tab[] = {
        {"UP", KB_KEY_UP}, {"DOWN", KB_KEY_DOWN}, {"LEFT", KB_KEY_LEFT}, {"RIGHT", KB_KEY_RIGHT}, {"W", KB_KEY_W},
        {"A", KB_KEY_A}, {"S", KB_KEY_S}, {"D", KB_KEY_D}, {"SPACE", KB_KEY_SPACE}, {"SHIFT", KB_KEY_LEFT_SHIFT}};
    for (size_t i = 0; i < sizeof tab / sizeof tab[0]; i++)
        if (!strcmp(s, tab[i].n)) return tab[i].k;
    return KB_KEY_UNKNOWN;
}

struct mfb_window* mfb_open_ex(const char* title, unsigned width, unsigned height, unsigned flags) {
    (void)title; (void)flags;
    struct mfb_window* w = (struct mfb_window*)calloc(1, sizeof *w);
    if (!w) return NULL;
    w->w = width; w->h = height;
    const char* f = getenv("RTB_FRAMES");
    w->frames_wanted = f ? atoi(f) : 240;
    if (w->frames_wanted < 1) w->frames_wanted = 1;
    w->stamp = (double*)calloc((size_t)w->frames_wanted + 1, sizeof(double));
    const char* k = getenv("RTB_KEYS");
    char buf[256];
    snprintf(buf, sizeof buf, "%s", k ? k : "RIGHT,D");
    for (char* tok = strtok(buf, ","); tok && w->nkeys < MAX_KEYS; tok = strtok(NULL, ",")) {
        mfb_key kk = key_by_name(tok);
        if (kk != KB_KEY_UNKNOWN) w->keys[w->nkeys++] = kk;
    }
    g_win = w;
    return w;
}
void mfb_set_keyboard_callback(struct mfb_window* w, mfb_keyboard_func cb) { if (w) w->on_key = cb; }

bool mfb_wait_sync(struct mfb_window* w) {
    if (!w) return false;
    if (w->frames_done == 0 && w->t_first_wait == 0.0) w->t_first_wait = now_s();
    if (w->frames_done >= w->frames_wanted) return false;
    if (w->on_key)
        for (int i = 0; i < w->nkeys; i++) w->on_key(w, w->keys[i], (mfb_key_mod)0, true);
    return true;
}
mfb_update_state mfb_update_ex(struct mfb_window* w, void* buffer, unsigned width, unsigned height) {
    if (!w) return STATE_INVALID_WINDOW;
    if (!buffer) return STATE_INVALID_BUFFER;
    w->w = width; w->h = height;
    w->last = buffer;
    if (w->frames_done < w->frames_wanted) w->stamp[w->frames_done] = now_s();
    w->frames_done++;
    return STATE_OK;
}
struct mfb_timer* mfb_timer_create(void) {
    struct mfb_timer* t = (struct mfb_timer*)calloc(1, sizeof *t);
    if (t) t->t0 = now_s();
    return t;
}
static int cmp_d(const void* a, const void* b) { double x = *(const double*)a, y = *(const double*)b; return (x > y) - (x < y); }
void mfb_timer_destroy(struct mfb_timer* t) {
    free(t);
    struct mfb_window* w = g_win;
    if (!w || w->frames_done == 0) return;
    int n = w->frames_done < w->frames_wanted ? w->frames_done : w->frames_wanted;
    double* dt = (double*)malloc(sizeof(double) * (size_t)n);
    double prev = w->t_first_wait;
    for (int i = 0; i < n; i++) { dt[i] = (w->stamp[i] - prev) * 1e3; prev = w->stamp[i]; }
    double total = w->stamp[n - 1] - w->t_first_wait;
    qsort(dt, (size_t)n, sizeof(double), cmp_d);
    printf("headless MiniFB: %d frames %ux%u, %.3f s, %.1f frames/s, frame time ms p50 %.3f p99 %.3f max %.3f\n", n, w->w,
           w->h, total, n / total, dt[n / 2], dt[(int)(0.99 * (n - 1))], dt[n - 1]);
    free(dt);
    const char* path = getenv("RTB_LAST_FRAME");
    if (path && w->last) {
        const uint32_t* px = (const uint32_t*)w->last;
        size_t np = (size_t)w->w * w->h;
        uint8_t* rgb = (uint8_t*)malloc(np * 3);
        for (size_t i = 0; i < np; i++) {
            rgb[3 * i] = (uint8_t)(px[i] >> 16); rgb[3 * i + 1] = (uint8_t)(px[i] >> 8); rgb[3 * i + 2] = (uint8_t)px[i];
        }
        rtpng_write_rgb8(path, rgb, w->w, w->h);
        free(rgb);
    }
}
