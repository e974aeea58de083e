// parsim_dump.inl -- result read-back and test dumps (included by parsim_step.inl).
namespace parsim {

// Particle 0 of THIS strip (alive array first, then the graveyard) and this strip's collisions.
int rank_result_local(Rank *c, double *x0, double *y0, int *found, long long *collisions) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (rank_ensure_sorted(c)) return 1;
    PS_CUDA(c, cudaMemsetAsync(c->d_found, 0, sizeof(int), c->stream));
    if (c->n_alive > 0)
        k_find_id<<<grid_for(c->n_alive), kThreads, 0, c->stream>>>(c->n_alive, c->id_cur, c->pos_latest,
                                                                    0u, c->d_p0, c->d_found);
    if (c->n_grave > 0)
        k_find_id<<<grid_for(c->n_grave), kThreads, 0, c->stream>>>(c->n_grave, c->grave_id,
                                                                    c->grave_pos, 0u, c->d_p0, c->d_found);
    if (check_launch(c, "k_find_id")) return 1;
    double2 p = make_double2(0, 0);
    int f = 0;
    PS_CUDA(c, cudaMemcpyAsync(&p, c->d_p0, sizeof(p), cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaMemcpyAsync(&f, c->d_found, sizeof(f), cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    *x0 = f ? p.x : 0.0;
    *y0 = f ? p.y : 0.0;
    *found = f;
    *collisions = (long long)c->h_ctr->collisions;
    return 0;
}

template <typename T>
static int fetch(Rank *c, T *host_dst, const T *dev, size_t n, const unsigned char *owned_h, bool merge) {
    if (!host_dst) return 0;
    if (!merge) {
        PS_CUDA(c, cudaMemcpy(host_dst, dev, sizeof(T) * n, cudaMemcpyDeviceToHost));
        return 0;
    }
    std::vector<T> tmp(n);
    PS_CUDA(c, cudaMemcpy(tmp.data(), dev, sizeof(T) * n, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) if (owned_h[i]) host_dst[i] = tmp[i];
    return 0;
}

int rank_dump(Rank *c, double *x, double *y, double *vx, double *vy, double *m, int *alive, int *cell,
              int *list_cell, unsigned long long *list_rank, unsigned char *owned) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (rank_ensure_sorted(c)) return 1;
    const size_t n = (size_t)c->n_part;
    double *dx, *dy, *dvx, *dvy, *dm;
    int *dalive, *dcell, *dlist;
    unsigned long long *drank;
    unsigned char *downed;
    PS_ALLOC(c, dx, n, true); PS_ALLOC(c, dy, n, true); PS_ALLOC(c, dvx, n, true);
    PS_ALLOC(c, dvy, n, true); PS_ALLOC(c, dm, n, true); PS_ALLOC(c, dalive, n, true);
    PS_ALLOC(c, dcell, n, true); PS_ALLOC(c, dlist, n, true); PS_ALLOC(c, drank, n, true);
    PS_ALLOC(c, downed, n, true);
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    const int n_tot = c->n_alive;  // guests are dumped by the strip they came from
    if (n_tot > 0)
        k_dump_alive<<<grid_for(n_tot), kThreads, 0, c->stream>>>(
            c->g, n_tot, c->pos_latest, c->vel_cur, c->m_cur, c->id_cur, c->cell_cur, c->newcell,
            c->flags, c->have_newcell, c->have_flags, dx, dy, dvx, dvy, dm, dalive, dcell, dlist, drank,
            downed);
    if (c->n_grave > 0)
        k_dump_dead<<<grid_for(c->n_grave), kThreads, 0, c->stream>>>(
            c->n_grave, c->grave_pos, c->grave_vel, c->grave_id, c->grave_cell, dx, dy, dvx, dvy, dm,
            dalive, dcell, dlist, drank, downed);
    if (check_launch(c, "k_dump")) return 1;
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    const bool merge = c->nranks > 1;
    std::vector<unsigned char> owned_h;
    if (merge || owned) {
        owned_h.resize(n);
        PS_CUDA(c, cudaMemcpy(owned_h.data(), downed, n, cudaMemcpyDeviceToHost));
        if (owned) memcpy(owned, owned_h.data(), n);
    }
    int rc = fetch(c, x, dx, n, owned_h.data(), merge) || fetch(c, y, dy, n, owned_h.data(), merge) ||
             fetch(c, vx, dvx, n, owned_h.data(), merge) || fetch(c, vy, dvy, n, owned_h.data(), merge) ||
             fetch(c, m, dm, n, owned_h.data(), merge) || fetch(c, alive, dalive, n, owned_h.data(), merge) ||
             fetch(c, cell, dcell, n, owned_h.data(), merge) ||
             fetch(c, list_cell, dlist, n, owned_h.data(), merge) ||
             fetch(c, list_rank, drank, n, owned_h.data(), merge);
    cudaFree(dx); cudaFree(dy); cudaFree(dvx); cudaFree(dvy); cudaFree(dm); cudaFree(dalive);
    cudaFree(dcell); cudaFree(dlist); cudaFree(drank); cudaFree(downed);
    return rc;
}

int rank_dump_cells(Rank *c, double *mx, double *my, double *m) {
    PS_CUDA(c, cudaSetDevice(c->device));
    if (c->nranks != 1) return c->fail("parsim_dump_cells: single-rank contexts only");
    const size_t C = (size_t)c->g.ncells;
    PS_CUDA(c, cudaStreamSynchronize(c->stream));
    std::vector<double2> xy(C);
    PS_CUDA(c, cudaMemcpy(xy.data(), c->com_xy + c->g.nc, sizeof(double2) * C, cudaMemcpyDeviceToHost));
    for (size_t k = 0; k < C; k++) {
        if (mx) mx[k] = xy[k].x;
        if (my) my[k] = xy[k].y;
    }
    if (m) PS_CUDA(c, cudaMemcpy(m, c->com_m + c->g.nc, sizeof(double) * C, cudaMemcpyDeviceToHost));
    return 0;
}

}  // namespace parsim
