// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// Phase-level access to the UNMODIFIED reference serial simulator.  The reference source is
// compiled from where it lies (-I/root/reference/serial, see oracle/Makefile); nothing is copied
// into this repo.  `private` is re-defined so that the test driver can call the four phase
// functions of ParticleSimulation one by one (serial/parsim.cpp:261-420) and read
// `particles`, `cells`, `cellParticles` (serial/parsim.cpp:200-202) after each of them.
//
// Exposed as a tiny C ABI (ctypes) so tests can pin oracle/parsim_oracle.c phase by phase
// against the real reference.  Built only where /root/reference exists (this container);
// the resulting oracle/_ref/libparsim_ref_phases.so travels to the GPU box prebuilt.
#include <cmath>
#include <iostream>
#include <vector>
#include <random>
#include <stdexcept>
#include <string>
#include <unordered_set>
#include <iomanip>
#include <algorithm>
#include <omp.h>

#define private public
#define main parsim_reference_main
#include "parsim.cpp"   // resolved through -I<reference>/serial
#undef main
#undef private

extern "C" {

void *refph_create(long seed, double side, long ncside, long long n_part) {
    return new ParticleSimulation(seed, side, ncside, n_part);
}
void refph_destroy(void *h) { delete static_cast<ParticleSimulation *>(h); }

// phase: 0 updateCOM, 1 updateForces, 2 applyForce loop only, 3 updateCellParticles,
//        4 checkCollisions   (serial/parsim.cpp:435-441, :379-386)
void refph_phase(void *h, int phase) {
    ParticleSimulation *s = static_cast<ParticleSimulation *>(h);
    switch (phase) {
    case 0: s->updateCOM(); break;
    case 1: s->updateForces(); break;
    case 2:
        for (size_t i = 0; i < s->particles.size(); i++)
            s->particles[i].applyForce(s->side_length, s->grid_size, s->cells);
        break;
    case 3: s->updateCellParticles(); break;
    case 4: s->checkCollisions(); break;
    }
}
void refph_step(void *h) {
    for (int p = 0; p < 5; p++) refph_phase(h, p);
}
long long refph_collisions(void *h) { return static_cast<ParticleSimulation *>(h)->collisions; }

// id-ordered particle state; any pointer may be NULL
void refph_particles(void *h, double *x, double *y, double *vx, double *vy, double *m,
                     double *fx, double *fy, int *alive, int *cell_index) {
    ParticleSimulation *s = static_cast<ParticleSimulation *>(h);
    for (size_t i = 0; i < s->particles.size(); i++) {
        const Particle &p = s->particles[i];
        if (x) x[i] = p.x;
        if (y) y[i] = p.y;
        if (vx) vx[i] = p.vx;
        if (vy) vy[i] = p.vy;
        if (m) m[i] = p.m;
        if (fx) fx[i] = p.fx;
        if (fy) fy[i] = p.fy;
        if (alive) alive[i] = p.alive ? 1 : 0;
        if (cell_index) cell_index[i] = p.cell_index;
    }
}
void refph_cells(void *h, double *mx, double *my, double *m) {
    ParticleSimulation *s = static_cast<ParticleSimulation *>(h);
    for (size_t c = 0; c < s->cells.size(); c++) {
        if (mx) mx[c] = s->cells[c].mx;
        if (my) my[c] = s->cells[c].my;
        if (m) m[c] = s->cells[c].m;
    }
}
// CSR dump of cellParticles: start[C+1], ids[total]; returns total (call with ids==NULL to size)
long long refph_lists(void *h, long long *start, long long *ids) {
    ParticleSimulation *s = static_cast<ParticleSimulation *>(h);
    long long tot = 0;
    for (size_t c = 0; c < s->cellParticles.size(); c++) {
        if (start) start[c] = tot;
        for (Particle *p : s->cellParticles[c]) {
            if (ids) ids[tot] = p - s->particles.data();
            tot++;
        }
    }
    if (start) start[s->cellParticles.size()] = tot;
    return tot;
}

}  // extern "C"
