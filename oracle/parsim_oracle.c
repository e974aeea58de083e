/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement ("oracle") of the reference serial simulator
 * MiguelParece/ParticleSimulation serial/parsim.cpp, in plain C.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may build, link or call this file;
 * the product library (particlesimulation_b200/csrc) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file (a) against the reference's
 * own golden vectors (serial/run_tests.sh:30-42, openMP/new_tests.sh:30-35), (b) bit for bit,
 * phase by phase, against the unmodified reference compiled from /root/reference
 * (oracle/ref_phases.cpp -> oracle/_ref/libparsim_ref_phases.so) and (c) against committed
 * fixtures generated from the reference serial binary (tests/golden/).
 *
 * Every function cites the reference lines it restates.  All arithmetic is IEEE double with
 * the reference's evaluation order; build with -O2 -ffp-contract=off (the reference is built
 * by g++ -O2 for baseline x86-64, i.e. without FMA contraction).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* serial/parsim.cpp:13-16 */
#define PO_G 6.67408e-11
#define PO_EPSILON2 (0.005 * 0.005)
#define PO_EPSILON 0.005
#define PO_DELTAT 0.1

/* ---- random numbers: serial/parsim.cpp:18-48 ------------------------------------------- */
typedef struct {
    uint32_t state;
    int use_normal;
} po_rng;

static void po_rng_seed(po_rng *r, int input_seed) {
    /* :24  seed(abs(input_seed) + 987654321), useNormal(input_seed < 0) */
    r->state = (uint32_t)abs(input_seed) + 987654321u;
    r->use_normal = input_seed < 0;
}

static double po_uniform01(po_rng *r) {
    /* :26-32  xorshift 13/17/5; the two states are added as 32-bit ints (wrapping) */
    uint32_t before = r->state;
    uint32_t s = before;
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
    r->state = s;
    int32_t sum = (int32_t)(before + s);
    return 0.5 + 0.2328306e-09 * sum;
}

static double po_normal01(po_rng *r) {
    /* :34-43  Box-Muller, mean 0.5, sigma 0.15, rejected outside [0,1) */
    double res;
    do {
        double u1 = po_uniform01(r);
        double u2 = po_uniform01(r);
        double z = sqrt(-2 * log(u1)) * cos(2 * M_PI * u2);
        res = 0.5 + 0.15 * z;
    } while (res < 0 || res >= 1);
    return res;
}

static double po_random01(po_rng *r) { /* :45-47 */
    return r->use_normal ? po_normal01(r) : po_uniform01(r);
}

/* serial/parsim.cpp:220-232 -- five draws per particle, strictly sequential */
void po_init_particles(long seed, double side, long ncside, long long n, double *x, double *y,
                       double *vx, double *vy, double *m) {
    po_rng rng;
    po_rng_seed(&rng, (int)seed); /* main stores the seed in an int, :465 */
    for (long long i = 0; i < n; i++) {
        x[i] = po_random01(&rng) * side;
        y[i] = po_random01(&rng) * side;
        vx[i] = (po_random01(&rng) - 0.5) * side / ncside / 5.0;
        vy[i] = (po_random01(&rng) - 0.5) * side / ncside / 5.0;
        m[i] = po_random01(&rng) * 0.01 * (ncside * ncside) / (size_t)n / PO_G * PO_EPSILON2;
    }
}

/* first `count` raw draws, for the known-answer tests */
void po_draws(long seed, int count, double *out) {
    po_rng rng;
    po_rng_seed(&rng, (int)seed);
    for (int i = 0; i < count; i++) out[i] = po_random01(&rng);
}

/* ---- simulation state ---------------------------------------------------------------- */
typedef struct {
    int *items;
    int len, cap;
} po_list;

typedef struct po_sim {
    long long n;
    long nc;
    long long ncells;
    double side;
    /* Particle, :52-80 (SoA here; id = index) */
    double *x, *y, *vx, *vy, *m, *fx, *fy;
    unsigned char *alive;
    int *cell_index;
    /* Cell, :82-107 */
    double *cmx, *cmy, *cm;
    int *cx, *cy;
    unsigned char *change_flag;
    /* cellParticles, :201 */
    po_list *lists;
    int semantics; /* 0 = serial/parsim.cpp (default), 1 = the OpenMP variant's list update */
    long long collisions;
    long long panics;
    /* scratch for checkCollisions */
    unsigned char *in_set;
    int *set_items;
    int set_cap;
} po_sim;

static void po_list_push(po_list *l, int v) {
    if (l->len == l->cap) {
        l->cap = l->cap ? 2 * l->cap : 8;
        l->items = (int *)realloc(l->items, sizeof(int) * (size_t)l->cap);
    }
    l->items[l->len++] = v;
}

po_sim *po_create_empty(double side, long ncside, long long n) {
    po_sim *s = (po_sim *)calloc(1, sizeof(po_sim));
    s->n = n;
    s->nc = ncside;
    s->ncells = (long long)ncside * ncside;
    s->side = side;
    size_t nn = (size_t)(n > 0 ? n : 1), cc = (size_t)(s->ncells > 0 ? s->ncells : 1);
    s->x = calloc(nn, 8); s->y = calloc(nn, 8); s->vx = calloc(nn, 8); s->vy = calloc(nn, 8);
    s->m = calloc(nn, 8); s->fx = calloc(nn, 8); s->fy = calloc(nn, 8);
    s->alive = malloc(nn); memset(s->alive, 1, nn);         /* :70 alive(true) */
    s->cell_index = calloc(nn, sizeof(int));
    s->cmx = calloc(cc, 8); s->cmy = calloc(cc, 8); s->cm = calloc(cc, 8);
    s->cx = calloc(cc, sizeof(int)); s->cy = calloc(cc, sizeof(int));
    s->change_flag = calloc(cc, 1);
    s->lists = calloc(cc, sizeof(po_list));
    s->in_set = calloc(nn, 1);
    return s;
}

/* ParticleSimulation ctor, :209-218 */
/* which variant's list update (phase 3) to restate: 0 = serial (the parity target), 1 = OpenMP */
void po_set_semantics(struct po_sim *s, int semantics);

po_sim *po_create(long seed, double side, long ncside, long long n) {
    po_sim *s = po_create_empty(side, ncside, n);
    po_init_particles(seed, side, ncside, n, s->x, s->y, s->vx, s->vy, s->m);
    return s;
}

po_sim *po_create_from(double side, long ncside, long long n, const double *x, const double *y,
                       const double *vx, const double *vy, const double *m) {
    po_sim *s = po_create_empty(side, ncside, n);
    memcpy(s->x, x, 8 * (size_t)n); memcpy(s->y, y, 8 * (size_t)n);
    memcpy(s->vx, vx, 8 * (size_t)n); memcpy(s->vy, vy, 8 * (size_t)n);
    memcpy(s->m, m, 8 * (size_t)n);
    return s;
}

void po_destroy(po_sim *s) {
    if (!s) return;
    for (long long c = 0; c < s->ncells; c++) free(s->lists[c].items);
    free(s->x); free(s->y); free(s->vx); free(s->vy); free(s->m); free(s->fx); free(s->fy);
    free(s->alive); free(s->cell_index); free(s->cmx); free(s->cmy); free(s->cm);
    free(s->cx); free(s->cy); free(s->change_flag); free(s->lists); free(s->in_set);
    free(s->set_items); free(s);
}

/* ---- phase 0: updateCOM, :261-290 (+ Cell::addParticle :93-106) ------------------------ */
static void po_update_com(po_sim *s) {
    for (long long c = 0; c < s->ncells; c++) { /* :263-264 both vectors re-assigned */
        s->lists[c].len = 0;
        s->cmx[c] = s->cmy[c] = s->cm[c] = 0;
        s->cx[c] = s->cy[c] = 0;
        s->change_flag[c] = 0;
    }
    const double cellw = s->side / s->nc;
    for (long long i = 0; i < s->n; i++) {
        int gx = (int)(s->x[i] / cellw); /* :268 */
        int gy = (int)(s->y[i] / cellw); /* :269 */
        int c = gy * (int)s->nc + gx;    /* :272 */
        if (gx < 0 || gx >= s->nc || gy < 0 || gy >= s->nc) { /* :276-280 */
            s->panics++;
            continue;
        }
        s->cell_index[i] = c;             /* :284 */
        po_list_push(&s->lists[c], (int)i); /* :285 */
        if (s->cm[c] == 0) {              /* :95-99 */
            s->cmx[c] = s->x[i];
            s->cmy[c] = s->y[i];
        } else {                          /* :102-103 */
            s->cmx[c] = (s->cmx[c] * s->cm[c] + s->m[i] * s->x[i]) / (s->cm[c] + s->m[i]);
            s->cmy[c] = (s->cmy[c] * s->cm[c] + s->m[i] * s->y[i]) / (s->cm[c] + s->m[i]);
        }
        s->cm[c] += s->m[i];              /* :105 */
        s->cx[c] = gx;                    /* :287-288 */
        s->cy[c] = gy;
    }
}

/* ---- phase 1: updateForces, :292-377 --------------------------------------------------- */
static inline void po_force_pair(po_sim *s, int a, int b) { /* :127-148 */
    double dx = s->x[b] - s->x[a];
    double dy = s->y[b] - s->y[a];
    double d2 = dx * dx + dy * dy;
    double d = sqrt(d2);
    if (d == 0) return;
    double f = (PO_G * s->m[a] * s->m[b]) / d2;
    double fxa = f * (dx / d);
    double fya = f * (dy / d);
    s->fx[a] += fxa; s->fy[a] += fya;
    s->fx[b] -= fxa; s->fy[b] -= fya;
}

static inline void po_force_cell(po_sim *s, int a, double mx, double my, double m) { /* :109-125 */
    double dx = mx - s->x[a];
    double dy = my - s->y[a];
    double d2 = dx * dx + dy * dy;
    double d = sqrt(d2);
    if (d == 0) return;
    double f = (PO_G * s->m[a] * m) / d2;
    s->fx[a] += f * (dx / d);
    s->fy[a] += f * (dy / d);
}

static void po_update_forces(po_sim *s) {
    const int nc = (int)s->nc;
    for (long long c = 0; c < s->ncells; c++) {
        double imx[8], imy[8], im[8];
        int ni = 0;
        for (int dx = -1; dx <= 1; dx++) {       /* :301 dx outer */
            for (int dy = -1; dy <= 1; dy++) {   /* :303 dy inner */
                if (dx == 0 && dy == 0) continue; /* :305 */
                int nx = s->cx[c] + dx, ny = s->cy[c] + dy;
                double sx = 0, sy = 0;
                if (nx >= nc) sx += s->side; else if (nx < 0) sx -= s->side; /* :314-321 */
                if (ny >= nc) sy += s->side; else if (ny < 0) sy -= s->side; /* :322-329 */
                nx = (nx + nc) % nc;              /* :332-333 */
                ny = (ny + nc) % nc;
                if (nx < 0 || nx >= nc || ny < 0 || ny >= nc) { s->panics++; continue; } /* :335 */
                int q = nx + ny * nc;             /* :341 */
                imx[ni] = sx + s->cmx[q];         /* :346-347 */
                imy[ni] = sy + s->cmy[q];
                im[ni] = s->cm[q];                /* :349 */
                ni++;
            }
        }
        po_list *l = &s->lists[c];
        for (int j = 0; j < l->len; j++) {        /* :356 */
            int a = l->items[j];
            for (int k = j + 1; k < l->len; k++) { /* :358-366 */
                int b = l->items[k];
                if (s->alive[a] && s->alive[b]) po_force_pair(s, a, b);
            }
            if (s->alive[a])                       /* :369-374 */
                for (int t = 0; t < ni; t++) po_force_cell(s, a, imx[t], imy[t], im[t]);
        }
    }
}

/* ---- phase 2: Particle::applyForce for every particle, :150-195, :54-60, :381-384 ------ */
static void po_apply_forces(po_sim *s) {
    const double cellw = s->side / s->nc;
    for (long long i = 0; i < s->n; i++) {
        if (s->m[i] == 0) { /* :151-155 dead particles never move */
            s->fx[i] = 0; s->fy[i] = 0;
            continue;
        }
        double ax = s->fx[i] / s->m[i];
        double ay = s->fy[i] / s->m[i];
        int gy = (int)(s->y[i] / cellw); /* :160-163 cell from the pre-move position */
        int gx = (int)(s->x[i] / cellw);
        s->cell_index[i] = gy * (int)s->nc + gx;
        s->x[i] += s->vx[i] * PO_DELTAT + 0.5 * ax * PO_DELTAT * PO_DELTAT; /* :55-56 */
        s->y[i] += s->vy[i] * PO_DELTAT + 0.5 * ay * PO_DELTAT * PO_DELTAT;
        s->vx[i] += ax * PO_DELTAT;       /* :58-59 */
        s->vy[i] += ay * PO_DELTAT;
        s->x[i] = fmod(s->x[i] + s->side, s->side); /* :171-172 */
        s->y[i] = fmod(s->y[i] + s->side, s->side);
        gy = (int)(s->y[i] / cellw);      /* :174-177 */
        gx = (int)(s->x[i] / cellw);
        int now = gy * (int)s->nc + gx;
        if (s->cell_index[i] != now) {    /* :183-190 */
            s->change_flag[s->cell_index[i]] = 1;
            s->cell_index[i] = now;
        }
        s->fx[i] = 0; s->fy[i] = 0;       /* :193-194 */
    }
}

/* ---- phase 3: updateCellParticles, :234-259 --------------------------------------------
 * The reference erases at the loop position and then still advances it (:251-253), so the
 * element that slides into the hole is never examined: of a run of consecutive leavers only
 * every other one is moved; the rest stay listed in their old cell for this step. */
void po_set_semantics(po_sim *s, int semantics) { s->semantics = semantics; }

static void po_update_cell_particles(po_sim *s) {
    for (long long c = 0; c < s->ncells; c++) {
        if (!s->change_flag[c]) continue;
        po_list *l = &s->lists[c];
        if (s->semantics == 1) {
            /* openMP/parsim-omp.cpp:248-278 with one thread: collect the leavers in list order,
             * append them to their new cells, then remove_if every leaver -- no strays */
            int w = 0;
            for (int k = 0; k < l->len; k++) {
                int p = l->items[k];
                if (s->cell_index[p] != c) po_list_push(&s->lists[s->cell_index[p]], p);
            }
            for (int k = 0; k < l->len; k++)
                if (s->cell_index[l->items[k]] == c) l->items[w++] = l->items[k];
            l->len = w;
            s->change_flag[c] = 0;
            continue;
        }
        for (int k = 0; k < l->len; k++) {
            int p = l->items[k];
            if (s->cell_index[p] != c) {
                po_list_push(&s->lists[s->cell_index[p]], p);
                memmove(&l->items[k], &l->items[k + 1], sizeof(int) * (size_t)(l->len - k - 1));
                l->len--;
            }
        }
        s->change_flag[c] = 0;
    }
}

/* ---- phase 4: checkCollisions, :388-420 (+ getDistance :76-79) ------------------------- */
static void po_check_collisions(po_sim *s) {
    for (long long c = 0; c < s->ncells; c++) {
        po_list *l = &s->lists[c];
        int nset = 0;
        for (int j = 0; j < l->len; j++) {
            int a = l->items[j];
            if (!s->alive[a]) continue;
            for (int k = j + 1; k < l->len; k++) {
                int b = l->items[k];
                if (!s->alive[b]) continue;
                double ddx = s->x[a] - s->x[b], ddy = s->y[a] - s->y[b];
                double dist = sqrt(ddx * ddx + ddy * ddy); /* pow(.,2) == square */
                if (dist < PO_EPSILON) {
                    if (!s->in_set[a] && !s->in_set[b]) s->collisions++; /* :404-405 */
                    if (nset + 2 > s->set_cap) {
                        s->set_cap = s->set_cap ? 2 * s->set_cap : 64;
                        s->set_items = realloc(s->set_items, sizeof(int) * (size_t)s->set_cap);
                    }
                    if (!s->in_set[a]) { s->in_set[a] = 1; s->set_items[nset++] = a; }
                    if (!s->in_set[b]) { s->in_set[b] = 1; s->set_items[nset++] = b; }
                }
            }
        }
        for (int t = 0; t < nset; t++) { /* :414-418 */
            int p = s->set_items[t];
            s->alive[p] = 0;
            s->m[p] = 0;
            s->in_set[p] = 0;
        }
    }
}

/* phase: 0 updateCOM, 1 updateForces, 2 applyForce loop, 3 updateCellParticles,
 *        4 checkCollisions -- same numbering as oracle/ref_phases.cpp */
void po_phase(po_sim *s, int phase) {
    switch (phase) {
    case 0: po_update_com(s); break;
    case 1: po_update_forces(s); break;
    case 2: po_apply_forces(s); break;
    case 3: po_update_cell_particles(s); break;
    case 4: po_check_collisions(s); break;
    }
}

void po_step(po_sim *s) { /* simulate body, :435-441 */
    for (int p = 0; p < 5; p++) po_phase(s, p);
}

void po_run(po_sim *s, long steps) { /* :432 */
    for (long t = 0; t < steps; t++) po_step(s);
}

long long po_collisions(const po_sim *s) { return s->collisions; }
long long po_panics(const po_sim *s) { return s->panics; }

void po_particles(const po_sim *s, double *x, double *y, double *vx, double *vy, double *m,
                  double *fx, double *fy, int *alive, int *cell_index) {
    size_t b = 8 * (size_t)s->n;
    if (x) memcpy(x, s->x, b);
    if (y) memcpy(y, s->y, b);
    if (vx) memcpy(vx, s->vx, b);
    if (vy) memcpy(vy, s->vy, b);
    if (m) memcpy(m, s->m, b);
    if (fx) memcpy(fx, s->fx, b);
    if (fy) memcpy(fy, s->fy, b);
    if (alive) for (long long i = 0; i < s->n; i++) alive[i] = s->alive[i];
    if (cell_index) memcpy(cell_index, s->cell_index, sizeof(int) * (size_t)s->n);
}

void po_cells(const po_sim *s, double *mx, double *my, double *m) {
    size_t b = 8 * (size_t)s->ncells;
    if (mx) memcpy(mx, s->cmx, b);
    if (my) memcpy(my, s->cmy, b);
    if (m) memcpy(m, s->cm, b);
}

long long po_lists(const po_sim *s, long long *start, long long *ids) {
    long long tot = 0;
    for (long long c = 0; c < s->ncells; c++) {
        if (start) start[c] = tot;
        for (int k = 0; k < s->lists[c].len; k++) {
            if (ids) ids[tot] = s->lists[c].items[k];
            tot++;
        }
    }
    if (start) start[s->ncells] = tot;
    return tot;
}

/* exact pair counts of the two O(n_c^2) loops for the current lists (roofline flops,
 * SURVEY 8(d)): force = alive pairs, collision = alive pairs of the post-move lists */
long long po_alive_pairs(const po_sim *s) {
    long long tot = 0;
    for (long long c = 0; c < s->ncells; c++) {
        long long a = 0;
        for (int k = 0; k < s->lists[c].len; k++) a += s->alive[s->lists[c].items[k]];
        tot += a * (a - 1) / 2;
    }
    return tot;
}

#ifdef PO_MAIN
/* same CLI and two-line stdout as serial/parsim.cpp:457-489, :450-453 */
#include <time.h>
int main(int argc, char **argv) {
    if (argc != 6) {
        fprintf(stderr, "Error: Usage: %s <seed> <side_length> <grid_size> <n_particles> <n_timesteps>\n",
                argv[0]);
        return 1;
    }
    long seed = strtol(argv[1], NULL, 10);
    double side = strtod(argv[2], NULL);
    long nc = strtol(argv[3], NULL, 10);
    long long n = strtoll(argv[4], NULL, 10);
    long steps = strtol(argv[5], NULL, 10);
    po_sim *s = po_create(seed, side, nc, n);
    const char *sem = getenv("PARSIM_SEMANTICS");
    if (sem && strcmp(sem, "omp") == 0) po_set_semantics(s, 1);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    po_run(s, steps);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    fprintf(stderr, "%.1fs\n", (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec));
    printf("%.3f %.3f\n%lld\n", s->x[0], s->y[0], s->collisions);
    po_destroy(s);
    return 0;
}
#endif
