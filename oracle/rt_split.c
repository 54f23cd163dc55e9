/* oracle/rt_split.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Proof-by-construction that the reference's per-pixel stack machine (raytracing.cl:28-195, restated in
 * rt_oracle.c:trace_pixel) splits into three phases with BIT-IDENTICAL output, which is what a kernel needs in order
 * to spread ONE heavy pixel (deep glass recursion: up to 2^15 segments) over many lanes / warps / CTAs:
 *
 *   phase 1  geometry walk  -- the DFS of raytracing.cl:41-190 without the light loop (:87-136).  Nothing in the
 *            walk depends on the RNG or on accumulated colour: hits, reflect/refract directions, Fresnel factors
 *            and the stack discipline are functions of the ray alone.  It emits, in program order, an event list:
 *            ADD(term) for a light hit (:52) or a skybox miss (:79), SHADE(k) for the k-th shaded segment,
 *            PUSH when a refraction child is stacked (:177), POP when an entry is folded into its parent (:188).
 *            This is 1/(1+2L) of the intersection tests.
 *   phase 2  shading        -- every SHADE record is independent: the ambient term (:83-84) and the 2L specular /
 *            diffuse terms (:129-135) need only (hit point, normal, material, f, segment origin) and the RNG state
 *            at the segment.  Each shaded segment draws EXACTLY 4L numbers (2 samples x 2 draws per light, no early
 *            exit in :87-136), so the state before segment k is xorshift32^(4Lk)(pixel id), and xorshift32
 *            (primitives.cl:116-125) is linear over GF(2): jump-ahead is a product of precomputed 32x32 bit
 *            matrices T^(2^j), T = step^(4L).  Phase 2 below visits the records in REVERSE order on purpose.
 *   phase 3  ordered replay -- additions only, in the event order, on a stack of partial sums: reproduces the
 *            reference's addition order exactly; 1+2L float3 adds per shaded segment.
 *
 * tests/test_split_cpu.py checks rto_split_render == rto_render word for word (render.map, glass-heavy grids, fuzz
 * scenes).  Built as librt_split.so by oracle/Makefile; it includes rt_oracle.c to reuse its static helpers. */
#include "rt_oracle.c"

/* ---------- xorshift32 as a GF(2) matrix: col[i] = image of bit i ---------- */
typedef struct { uint32_t col[32]; } bitmat;

static inline uint32_t xs_step(uint32_t x) { x ^= x << 13; x ^= x >> 17; x ^= x << 5; return x; } /* primitives.cl:119-121 */

static inline uint32_t bm_apply(const bitmat* m, uint32_t x) {
    uint32_t r = 0;
    for (int i = 0; i < 32; i++) if (x >> i & 1u) r ^= m->col[i];
    return r;
}
static void bm_mul(bitmat* out, const bitmat* a, const bitmat* b) { /* out = a after b */
    bitmat t;
    for (int i = 0; i < 32; i++) t.col[i] = bm_apply(a, b->col[i]);
    *out = t;
}
typedef struct { bitmat pw[32]; } jump_table;   /* pw[j] = (step^draws)^(2^j) */

static void jump_init(jump_table* jt, uint32_t draws_per_segment) {
    bitmat one, acc;
    for (int i = 0; i < 32; i++) { one.col[i] = xs_step(1u << i); acc.col[i] = 1u << i; }
    for (uint32_t d = 0; d < draws_per_segment; d++) bm_mul(&acc, &one, &acc);
    jt->pw[0] = acc;
    for (int j = 1; j < 32; j++) bm_mul(&jt->pw[j], &jt->pw[j - 1], &jt->pw[j - 1]);
}
static uint32_t jump_to_segment(const jump_table* jt, uint32_t seed, uint32_t k) {
    for (int j = 0; k; j++, k >>= 1) if (k & 1u) seed = bm_apply(&jt->pw[j], seed);
    return seed;
}
uint32_t rto_split_jump(uint32_t seed, uint32_t draws_per_segment, uint32_t k) { /* KAT entry point */
    jump_table jt; jump_init(&jt, draws_per_segment);
    return jump_to_segment(&jt, seed, k);
}

/* ---------- phase 1: geometry walk ---------- */
enum { EV_ADD = 0, EV_SHADE = 1, EV_PUSH = 2, EV_POP = 3 };
typedef struct { uint32_t kind; uint32_t k; v3 term; } event_t;
typedef struct { v3 ip, nrm, from; float f; w_material m; v3 terms[1 + 2 * 255]; uint32_t nterms; } shade_rec;

typedef struct {
    event_t* ev; size_t nev, cap_ev;
    shade_rec* rec; size_t nrec, cap_rec;
} pixel_plan;

static event_t* plan_event(pixel_plan* p) {
    if (p->nev == p->cap_ev) { p->cap_ev = p->cap_ev ? 2 * p->cap_ev : 64; p->ev = realloc(p->ev, p->cap_ev * sizeof(event_t)); }
    return &p->ev[p->nev++];
}
static shade_rec* plan_record(pixel_plan* p) {
    if (p->nrec == p->cap_rec) { p->cap_rec = p->cap_rec ? 2 * p->cap_rec : 16; p->rec = realloc(p->rec, p->cap_rec * sizeof(shade_rec)); }
    return &p->rec[p->nrec++];
}

/* Same control flow as rt_oracle.c:trace_pixel (raytracing.cl:41-190) with the light loop (:87-136) removed. */
static void walk_pixel(const float* cam, uint32_t W, uint32_t id, const scene_t* sc, pixel_plan* plan) {
    typedef struct { v3 o, d; int depth; float n, f; } walk_entry;
    walk_entry st[STACK_MAX];
    uint32_t sz = 1;
    plan->nev = plan->nrec = 0;
    primary_ray(cam, W, id, &st[0].o, &st[0].d);
    st[0].depth = 0; st[0].n = AIR_N; st[0].f = 1.0f;
    while (sz > 0) {
        while (st[sz - 1].depth < STACK_MAX) {
            walk_entry* top = &st[sz - 1];
            v3 lcol;
            if (find_light(top->o, top->d, sc, &lcol, 0)) {                    /* :48-54 */
                event_t* e = plan_event(plan); e->kind = EV_ADD; e->k = 0; e->term = scl(lcol, top->f);
                break;
            }
            v3 ip, nrm; w_material m;
            if (!find_solid(top->o, top->d, sc, &ip, &nrm, &m, 0)) {           /* :56-81 */
                int u, v;
                cube_uv(top->d, (uint32_t)(sc->sw / 4), &u, &v);
                v3 px = texel(sc->sky, sc->sw, sc->sh, 1, u, sc->sh - v, 0);
                event_t* e = plan_event(plan); e->kind = EV_ADD; e->k = 0; e->term = scl(px, top->f);
                break;
            }
            {   /* the shaded segment: everything :83-136 reads, captured BEFORE :150-159 update the entry */
                uint32_t k = (uint32_t)plan->nrec;
                shade_rec* r = plan_record(plan);
                r->ip = ip; r->nrm = nrm; r->from = top->o; r->f = top->f; r->m = m; r->nterms = 0;
                event_t* e = plan_event(plan); e->kind = EV_SHADE; e->k = k; e->term = mk(0, 0, 0);
            }
            v3 incident = top->d;
            float n1 = top->n;
            float n2 = (n1 == AIR_N) ? m.n : AIR_N;
            float refl = m.reflectivity;
            if (m.dielectric) {
                float fr = fresnel_schlick(n1, n2, incident, nrm);
                refl = m.reflectivity + (1.0f - m.reflectivity) * fr;
            }
            float old_f = top->f;
            top->f *= refl;
            top->d = mirror_dir(top->d, nrm);
            top->o = ip;
            top->depth++;
            if (m.transperent && sz < STACK_MAX && refl < 1.0f) {
                walk_entry* ch = &st[sz];
                *ch = *top;
                if (n1 < n2) ch->o = sub(ch->o, scl(nrm, 2.0f * EPS_OFFSET));
                else nrm = scl(nrm, -1.0f);
                ch->f = old_f * (1.0f - refl);
                ch->n = n2;
                if (!bend_dir(n1, n2, incident, nrm, &ch->d) || ch->d.x != ch->d.x) continue;
                sz++;
                plan_event(plan)->kind = EV_PUSH;
            }
        }
        if (sz == 1) break;
        plan_event(plan)->kind = EV_POP;                                       /* :188 */
        sz--;
    }
}

/* ---------- phase 2: one shaded segment, from its record and the jumped RNG state alone (:83-136) ---------- */
static void shade_segment(shade_rec* r, uint32_t rng, const scene_t* sc) {
    uint32_t n = 0;
    r->terms[n++] = scl(scl(ld3(r->m.rgb), r->f), r->m.ambient);
    for (uint32_t i = 0; i < sc->L; i++) {
        const w_light* li = &sc->lit[i];
        v3 lo = ld3(li->origin);
        float soft = 0.0f;
        v3 sdir = unit(sub(lo, r->ip));
        for (int j = 0; j < SHADOW_SAMPLES; j++) {
            float theta = (float)(2 * M_PI * (double)rng_next(&rng));
            float phi = (float)(M_PI * (double)rng_next(&rng));
            float x = li->radius * sinf(phi) * cosf(theta);
            float y = li->radius * sinf(phi) * sinf(theta);
            float z = li->radius * cosf(phi);
            soft += shadow_path(add(lo, mk(x, y, z)), r->ip, sc);
        }
        float ssr = soft / (float)SHADOW_SAMPLES;
        float d = dist3(lo, r->ip);
        v3 lrgb = dvs(scl(scl(scl(ld3(li->rgb), li->intensity), INV_PI_F), 1.0f), d * d);
        lrgb = scl(lrgb, ssr);
        v3 vv = unit(sub(r->from, r->ip));
        v3 hh = unit(add(vv, sdir));
        float ndh = dot3(r->nrm, hh);
        float spec = powf(0.0f < ndh ? ndh : 0.0f, (float)r->m.shininess);
        r->terms[n++] = scl(scl(lrgb, r->f * r->m.specular), spec);
        float ndl = dot3(r->nrm, sdir);
        r->terms[n++] = scl(scl(lrgb, r->f * r->m.diffuse), 0.0f < ndl ? ndl : 0.0f);
    }
    r->nterms = n;
}

/* ---------- phase 3: ordered replay (additions only) ---------- */
static uint32_t replay_pixel(const pixel_plan* plan) {
    v3 acc[STACK_MAX];
    uint32_t sz = 1;
    acc[0] = mk(0, 0, 0);
    for (size_t i = 0; i < plan->nev; i++) {
        const event_t* e = &plan->ev[i];
        switch (e->kind) {
        case EV_ADD: acc[sz - 1] = add(acc[sz - 1], e->term); break;
        case EV_SHADE: {
            const shade_rec* r = &plan->rec[e->k];
            for (uint32_t t = 0; t < r->nterms; t++) acc[sz - 1] = add(acc[sz - 1], r->terms[t]);
            break;
        }
        case EV_PUSH: acc[sz++] = mk(0, 0, 0); break;
        default: acc[sz - 2] = add(acc[sz - 2], acc[sz - 1]); sz--; break;
        }
    }
    v3 c = acc[0];
    float r = CLAMP01(c.x) * 255.0f, g = CLAMP01(c.y) * 255.0f, b = CLAMP01(c.z) * 255.0f;
    return ((uint32_t)r << 16) | ((uint32_t)g << 8) | (uint32_t)b;
}

/* Rows [y0,y1) through the three phases; same arguments as rto_render (single-threaded; triangles from
 * rto_set_triangles of THIS library).  shaded (may be NULL) receives the number of shaded segments per pixel. */
int rto_split_render(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                     const void* spheres, uint32_t S, const void* planes, uint32_t P, const void* lights, uint32_t L,
                     const uint8_t* tex, int tw, int th, int tl, const uint8_t* sky, int sw, int sh, uint32_t* out,
                     uint32_t* shaded) {
    if (!wide && (S > 255 || P > 255 || L > 255)) return -1;
    if (y1 > H || y0 > y1 || W == 0 || L > 255) return -1;
    scene_t sc = {(const w_sphere*)spheres, S, (const w_plane*)planes, P, (const w_light*)lights, L,
                  tex, tw, th, tl, sky, sw, sh, g_tri, g_T};
    jump_table jt;
    jump_init(&jt, 2u * SHADOW_SAMPLES * L);      /* draws per shaded segment: 2 per sample, :99-100 */
    pixel_plan plan = {0};
    for (uint32_t id = y0 * W; id < y1 * W; id++) {
        walk_pixel(cam, W, id, &sc, &plan);
        for (size_t k = plan.nrec; k-- > 0;)      /* reverse order on purpose: the records are independent */
            shade_segment(&plan.rec[k], jump_to_segment(&jt, id, (uint32_t)k), &sc);
        out[id] = replay_pixel(&plan);
        if (shaded) shaded[id] = (uint32_t)plan.nrec;
    }
    free(plan.ev); free(plan.rec);
    return 0;
}

/* =====================================================================================================================
 * Tree form: the same pixel with its segments evaluated in BREADTH-FIRST order.
 *
 * A shaded segment has up to two continuations whose geometry is independent of one another: the refraction child
 * (a new stack entry, raytracing.cl:161-179) and the reflection continuation of its own entry (:150-159).  Both are
 * functions of the segment's ray state {origin, direction, depth, n, f, stack height} alone, so the segments of a
 * pixel form a binary tree that can be expanded level by level -- on a GPU: one frontier node per lane -- instead of
 * one after the other.  The stack machine's order is the tree's pre-order with the refraction sub-tree first; the
 * k of "k-th shaded segment" (hence the RNG state, 4L draws each) is the pre-order rank among shaded nodes, and the
 * colour is the pre-order list of additions with PUSH before / POP after every refraction entry (an entry pushed at
 * depth 15 has no segment but is still pushed and popped: parent += 0).  rto_tree_render builds the tree from a FIFO
 * queue, ranks it, shades the nodes from the LAST to the first, replays -- and must equal rto_render word for word.
 * ===================================================================================================================== */
enum { ND_LIGHT = 0, ND_SKY = 1, ND_SHADED = 2 };
typedef struct {
    v3 o, d; int depth; float n, f; uint32_t sz;   /* ray state on entry to the segment */
    int kind; v3 term;                              /* light / sky term, or the ambient term */
    int entry_child;                                /* a refraction entry was pushed after this segment */
    int32_t refr, refl;                             /* first segment of that entry / next segment of this entry (-1: none) */
    uint32_t rank;                                  /* pre-order rank among shaded nodes */
    shade_rec rec;
} tree_node;

typedef struct { tree_node* nd; size_t n, cap; } pixel_tree;

static int32_t tree_new(pixel_tree* t, v3 o, v3 d, int depth, float n, float f, uint32_t sz) {
    if (t->n == t->cap) { t->cap = t->cap ? 2 * t->cap : 32; t->nd = realloc(t->nd, t->cap * sizeof(tree_node)); }
    tree_node* x = &t->nd[t->n];
    x->o = o; x->d = d; x->depth = depth; x->n = n; x->f = f; x->sz = sz;
    x->entry_child = 0; x->refr = x->refl = -1; x->rank = 0; x->rec.nterms = 0;
    return (int32_t)t->n++;
}

/* One segment (raytracing.cl:42-179 without :87-136), from the node's own state; appends its continuations. */
static void tree_expand(pixel_tree* t, int32_t at, const scene_t* sc) {
    tree_node x = t->nd[at];            /* by value: tree_new may move the array */
    v3 lcol;
    if (find_light(x.o, x.d, sc, &lcol, 0)) { t->nd[at].kind = ND_LIGHT; t->nd[at].term = scl(lcol, x.f); return; }
    v3 ip, nrm; w_material m;
    if (!find_solid(x.o, x.d, sc, &ip, &nrm, &m, 0)) {
        int u, v;
        cube_uv(x.d, (uint32_t)(sc->sw / 4), &u, &v);
        t->nd[at].kind = ND_SKY;
        t->nd[at].term = scl(texel(sc->sky, sc->sw, sc->sh, 1, u, sc->sh - v, 0), x.f);
        return;
    }
    t->nd[at].kind = ND_SHADED;
    t->nd[at].term = scl(scl(ld3(m.rgb), x.f), m.ambient);
    t->nd[at].rec.ip = ip; t->nd[at].rec.nrm = nrm; t->nd[at].rec.from = x.o; t->nd[at].rec.f = x.f; t->nd[at].rec.m = m;
    float n1 = x.n, n2 = (n1 == AIR_N) ? m.n : AIR_N;
    float refl = m.reflectivity;
    if (m.dielectric) refl = m.reflectivity + (1.0f - m.reflectivity) * fresnel_schlick(n1, n2, x.d, nrm);
    float f_refl = x.f * refl;
    v3 d_refl = mirror_dir(x.d, nrm);
    int depth = x.depth + 1;
    if (m.transperent && x.sz < STACK_MAX && refl < 1.0f) {
        v3 co = ip, nn = nrm, cd;
        if (n1 < n2) co = sub(ip, scl(nrm, 2.0f * EPS_OFFSET));
        else nn = scl(nrm, -1.0f);
        if (bend_dir(n1, n2, x.d, nn, &cd) && !(cd.x != cd.x)) {
            t->nd[at].entry_child = 1;
            if (depth < STACK_MAX) {
                int32_t c = tree_new(t, co, cd, depth, n2, x.f * (1.0f - refl), x.sz + 1);
                t->nd[at].refr = c;
            }
        }
    }
    if (depth < STACK_MAX) {
        int32_t c = tree_new(t, ip, d_refl, depth, x.n, f_refl, x.sz);
        t->nd[at].refl = c;
    }
}

static void tree_rank(pixel_tree* t, int32_t at, uint32_t* next) {     /* pre-order: refraction entry first */
    while (at >= 0) {
        tree_node* x = &t->nd[at];
        if (x->kind != ND_SHADED) return;
        x->rank = (*next)++;
        if (x->refr >= 0) tree_rank(t, x->refr, next);
        at = t->nd[at].refl;
    }
}

static v3 tree_replay(const pixel_tree* t, int32_t at) {                 /* the colour of one stack entry */
    v3 acc = mk(0, 0, 0);
    while (at >= 0) {
        const tree_node* x = &t->nd[at];
        acc = add(acc, x->term);
        if (x->kind != ND_SHADED) break;
        for (uint32_t i = 0; i < x->rec.nterms; i++) acc = add(acc, x->rec.terms[i]);
        if (x->entry_child) acc = add(acc, tree_replay(t, x->refr));      /* PUSH ... POP: parent += child (0 if empty) */
        at = x->refl;
    }
    return acc;
}

/* The same two steps without recursion, the way a warp would do them over the flat, level-ordered node array
 * (children always have larger indices than their parent):
 *   ranks   shaded-node counts of every sub-tree by one pass from the last node to the first (a level at a time on a
 *           GPU), then ranks from the first node to the last: rank(refr child) = rank + 1,
 *           rank(refl child) = rank + 1 + count(refr sub-tree);
 *   replay  pre-order with an explicit stack of {node to resume at, partial sum}: <= 15 frames (raytracing.cl:9). */
static void tree_rank_flat(pixel_tree* t, uint32_t* count /* t->n scratch */) {
    for (size_t i = t->n; i-- > 0;) {
        const tree_node* x = &t->nd[i];
        count[i] = (x->kind == ND_SHADED ? 1u : 0u) + (x->refr >= 0 ? count[x->refr] : 0u) + (x->refl >= 0 ? count[x->refl] : 0u);
    }
    t->nd[0].rank = 0;
    for (size_t i = 0; i < t->n; i++) {
        const tree_node* x = &t->nd[i];
        if (x->kind != ND_SHADED) continue;
        if (x->refr >= 0) t->nd[x->refr].rank = x->rank + 1u;
        if (x->refl >= 0) t->nd[x->refl].rank = x->rank + 1u + (x->refr >= 0 ? count[x->refr] : 0u);
    }
}
static v3 tree_replay_flat(const pixel_tree* t) {
    struct { int32_t resume; v3 acc; } fr[STACK_MAX];
    int sp = 0;
    int32_t cur = 0;
    v3 acc = mk(0, 0, 0);
    for (;;) {
        int ended = 1;
        if (cur >= 0) {
            const tree_node* x = &t->nd[cur];
            acc = add(acc, x->term);
            if (x->kind == ND_SHADED) {
                for (uint32_t i = 0; i < x->rec.nterms; i++) acc = add(acc, x->rec.terms[i]);
                if (x->entry_child) {                       /* PUSH */
                    fr[sp].resume = x->refl; fr[sp].acc = acc; sp++;
                    acc = mk(0, 0, 0);
                    cur = x->refr;                          /* -1: an entry without segments, ends at once */
                    ended = cur < 0;
                } else if (x->refl >= 0) { cur = x->refl; ended = 0; }
            }
        }
        while (ended) {                                     /* the entry is finished: POP into its parent and resume it */
            if (sp == 0) return acc;
            sp--;
            acc = add(fr[sp].acc, acc);
            cur = fr[sp].resume;
            ended = cur < 0;
        }
    }
}
static int g_tree_flat = 0;
void rto_tree_mode(int flat) { g_tree_flat = flat; }        /* 1: rank and replay without recursion */

int rto_tree_render(int wide, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                    const void* spheres, uint32_t S, const void* planes, uint32_t P, const void* lights, uint32_t L,
                    const uint8_t* tex, int tw, int th, int tl, const uint8_t* sky, int sw, int sh, uint32_t* out,
                    uint32_t* levels /* may be NULL: warp steps << 24 | levels << 16 | segments */) {
    if (!wide && (S > 255 || P > 255 || L > 255)) return -1;
    if (y1 > H || y0 > y1 || W == 0 || L > 255) return -1;
    scene_t sc = {(const w_sphere*)spheres, S, (const w_plane*)planes, P, (const w_light*)lights, L,
                  tex, tw, th, tl, sky, sw, sh, g_tri, g_T};
    jump_table jt;
    jump_init(&jt, 2u * SHADOW_SAMPLES * L);
    pixel_tree t = {0};
    uint32_t* count = 0; size_t ncount = 0;
    for (uint32_t id = y0 * W; id < y1 * W; id++) {
        t.n = 0;
        v3 o, d;
        primary_ray(cam, W, id, &o, &d);
        tree_new(&t, o, d, 0, AIR_N, 1.0f, 1);
        uint32_t nlev = 0, steps = 0;                         /* steps: passes of a 32-lane warp taking one level at a time */
        for (size_t head = 0; head < t.n;) {                 /* FIFO = level order; a level is what a warp would take at once */
            size_t level_end = t.n;
            steps += (uint32_t)((level_end - head + 31) / 32);
            for (; head < level_end; head++) tree_expand(&t, (int32_t)head, &sc);
            nlev++;
        }
        uint32_t nshaded = 0;
        if (g_tree_flat) {
            if (t.n > ncount) { ncount = 2 * t.n; count = realloc(count, ncount * sizeof(uint32_t)); }
            tree_rank_flat(&t, count);
        } else tree_rank(&t, 0, &nshaded);
        for (size_t k = t.n; k-- > 0;)                        /* last node first: the records are independent */
            if (t.nd[k].kind == ND_SHADED) {
                shade_rec* r = &t.nd[k].rec;
                shade_segment(r, jump_to_segment(&jt, id, t.nd[k].rank), &sc);
                for (uint32_t i = 1; i < r->nterms; i++) r->terms[i - 1] = r->terms[i];   /* the ambient term lives in node.term */
                r->nterms--;
            }
        v3 c = g_tree_flat ? tree_replay_flat(&t) : tree_replay(&t, 0);
        float r = CLAMP01(c.x) * 255.0f, g = CLAMP01(c.y) * 255.0f, b = CLAMP01(c.z) * 255.0f;
        out[id] = ((uint32_t)r << 16) | ((uint32_t)g << 8) | (uint32_t)b;
        if (levels) levels[id] = ((steps > 255 ? 255u : steps) << 24) | (nlev << 16) | (uint32_t)(t.n > 0xffff ? 0xffff : t.n);
    }
    free(t.nd); free(count);
    return 0;
}
