// grb_mxv.cu — GrB_mxv and GrB_vxm: argument checking, direction choice (push / streaming pull / early-exit
// pull), type reconciliation and the launch sequence.
// Replaces what python-graphblas reaches through `call("GrB_mxv"|"GrB_vxm", [w, mask, accum, op, A|u, u|A, desc])`
// (reference graphblas/core/base.py:476-483; Matrix.mxv core/matrix.py:2163-2204; Vector.vxm core/vector.py:1259-1306).
#include <cub/cub.cuh>

#include "grb_internal.cuh"
#include "grb_ops.cuh"
#include "grb_spmv.cuh"
#include "grb_spmv_launch.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

bool spmv_launch_pull(int mon, int mul, int tcode, const void *pa, int mode) {
    switch (mon) {
    case OP_PLUS: return spmv_pull_PLUS(mul, tcode, pa, mode);
    case OP_MIN: return spmv_pull_MIN(mul, tcode, pa, mode);
    case OP_MAX: return spmv_pull_MAX(mul, tcode, pa, mode);
    case OP_ANY: return spmv_pull_ANY(mul, tcode, pa, mode);
    case OP_LOR: case OP_LAND: case OP_LXOR: case OP_LXNOR: return spmv_pull_BOOLMON(mul | (mon << 8), tcode, pa, mode);
    }
    return false;
}
bool spmv_launch_stream(int mon, int mul, int tcode, const void *sa) {
    switch (mon) {
    case OP_PLUS: return spmv_stream_PLUS(mul, tcode, sa);
    case OP_MIN: return spmv_stream_MIN(mul, tcode, sa);
    case OP_MAX: return spmv_stream_MAX(mul, tcode, sa);
    case OP_ANY: return spmv_stream_ANY(mul, tcode, sa);
    case OP_LOR: case OP_LAND: case OP_LXOR: case OP_LXNOR: return spmv_stream_BOOLMON(mul | (mon << 8), tcode, sa);
    }
    return false;
}
bool spmv_launch_red(int mon, int mul, int tcode, const void *ra) {
    switch (mon) {
    case OP_PLUS: return spmv_red_PLUS(mul, tcode, ra);
    case OP_MIN: return spmv_red_MIN(mul, tcode, ra);
    case OP_MAX: return spmv_red_MAX(mul, tcode, ra);
    case OP_ANY: return spmv_red_ANY(mul, tcode, ra);
    case OP_LOR: case OP_LAND: case OP_LXOR: case OP_LXNOR: return spmv_red_BOOLMON(mul | (mon << 8), tcode, ra);
    }
    return false;
}
bool spmv_launch_push(int mon, int mul, int tcode, const void *pa) {
    switch (mon) {
    case OP_PLUS: return spmv_push_PLUS(mul, tcode, pa);
    case OP_MIN: return spmv_push_MIN(mul, tcode, pa);
    case OP_MAX: return spmv_push_MAX(mul, tcode, pa);
    case OP_ANY: return spmv_push_ANY(mul, tcode, pa);
    case OP_LOR: case OP_LAND: case OP_LXOR: case OP_LXNOR: return spmv_push_BOOLMON(mul | (mon << 8), tcode, pa);
    }
    return false;
}

int hub_capacity_for(size_t tsize) { return tsize > 4 ? hub_capacity<double>() : hub_capacity<float>(); }

template <class T> static T monoid_identity(int mon) {
    switch (mon) {
    case OP_MIN:
        if constexpr (std::is_same<T, bool>::value) return true;
        else if constexpr (std::is_floating_point<T>::value) return std::numeric_limits<T>::infinity();
        else return std::numeric_limits<T>::max();
    case OP_MAX:
        if constexpr (std::is_same<T, bool>::value) return false;
        else if constexpr (std::is_floating_point<T>::value) return -std::numeric_limits<T>::infinity();
        else return std::numeric_limits<T>::lowest();
    case OP_LAND: case OP_LXNOR: case OP_TIMES: return (T)1;
    default: return (T)0;
    }
}

DescFlags desc_flags(const GrB_Descriptor d) {
    DescFlags f = {false, false, false, false, false, 0};
    if (d) {
        if (d->magic != GRB_MAGIC) grb_throw(GrB_INVALID_OBJECT, "descriptor is not valid");
        f.replace = d->replace; f.comp = d->comp; f.structure = d->structure; f.tran0 = d->tran0; f.tran1 = d->tran1;
        f.axb = d->axb_method;
    }
    return f;
}

template <class T>
static EpiArgs<T> make_epi(GB_Object *w, bool pingpong, GB_Object *mask, const DescFlags &f, int accum) {
    EpiArgs<T> e;
    e.w_in = (const T *)w->vals; e.wp_in = w->present;
    if (pingpong) { vec_ensure_alt(w); e.w_out = (T *)w->vals_alt; e.wp_out = w->present_alt; }
    else { e.w_out = (T *)w->vals; e.wp_out = w->present; }
    e.mask_p = mask ? mask->present : nullptr;
    e.mask_v = mask ? mask->vals : nullptr;
    e.mask_vsize = (mask && !f.structure) ? (int)mask->type->size : 0;
    e.mask_fp = mask ? (mask->type->code == TC_FP32 || mask->type->code == TC_FP64) : 0;
    e.comp = f.comp; e.replace = f.replace; e.accum = accum;
    e.count = (unsigned long long *)w->d_nvals;
    return e;
}
static void swap_pingpong(GB_Object *w) {
    std::swap(w->vals, w->vals_alt);
    std::swap(w->present, w->present_alt);
}

// t = M (+).(x) u computed into (w or a temporary) with the requested write-back.
//   Mpull: CSR whose rows are indexed by the OUTPUT index (used by pull)
//   Mpush: CSR whose rows are indexed by u's index        (used by push); fetched lazily through getters
// one-shot request of GxB_b200_next_product_progress, taken over by the product that follows it
struct ProgressReq { unsigned long long *counters = nullptr, *expected = nullptr; int nblocks = 0; int64_t rpb = 0; };

template <class T>
static void product_typed(GB_Object *w, GB_Object *mask, int accum, int mon, int mul, int tcode, GB_Object *Aobj,
                          bool use_transpose, const Csr *Aoverride, GB_Object *u, const void *u_vals,
                          const DescFlags &f, int narrow = -1, ProgressReq prog = ProgressReq()) {
    // narrow >= 0: the semiring's type is that 8/16-bit integer type and T is its 32-bit stand-in (grb_internal.cuh):
    // results always pass through a temporary and are wrapped to the narrow type before the write-back
    const int64_t n_out = w->nrows, n_in = u->nrows;
    const bool fused = (w->type->code == tcode) && narrow < 0;
    const int64_t nu = vec_nvals(u);
    // ---- direction ----
    bool push = (nu * g.push_ratio < n_in);
    if (prog.counters) push = false;      // progress is reported by the tiled pull only
    if (g.force_direction == 1) push = false;
    if (g.force_direction == 2) push = true;
    if (nu == 0) push = true;   // nothing to multiply: the scatter is skipped, only the write-back runs
    if (prog.counters && push)
        grb_throw(GrB_NOT_IMPLEMENTED, "progress reporting (GxB_b200_next_product_progress) needs a non-empty, full input vector");
    const bool any_mon = (mon == OP_ANY);
    const bool have_mask = (mask != nullptr);

    // A's values in the semiring's type; multiplies that ignore them (second, pair) never touch the pointer
    const bool uses_x = !(mul == OP_SECOND || mul == OP_ONEB);
    auto vals_of = [&](Csr *M) -> const T * {
        return (const T *)(uses_x ? csr_vals_as(*M, Aobj->type->code, tcode, narrow) : M->vals);
    };
    auto get_csr = [&](bool want_transpose) -> Csr * {
        if (Aoverride) return nullptr;
        if (want_transpose) { mat_ensure_csc(Aobj); return &Aobj->csc; }
        mat_ensure_csr(Aobj);
        return &Aobj->csr;
    };

    // temporary for unfused / push paths
    T *t = nullptr; uint8_t *tp = nullptr;
    auto writeback = [&](const T *tv, const uint8_t *tpres) {
        // w<m,replace> = accum(w, t) in w's own type
        dzero(w->d_nvals, sizeof(int64_t));
        if (fused) {
            EpiArgs<T> e = make_epi<T>(w, false, mask, f, accum);
            launch_vec_writeback<T>(e, tv, tpres, n_out);
        } else {
            int wcode = w->type->code;
            void *tc = dmalloc((size_t)(n_out > 0 ? n_out : 1) * w->type->size);
            if (narrow >= 0 && wcode != narrow) {       // wrap to the semiring's own type first
                void *tn = dmalloc((size_t)(n_out > 0 ? n_out : 1) * kTypeSize[narrow]);
                cast_array(tv, tcode, tn, narrow, n_out);
                cast_array(tn, narrow, tc, wcode, n_out);
                dfree(tn);
            } else cast_array(tv, tcode, tc, wcode, n_out);
            GRB_TYPE_SWITCH(wcode, W, {
                EpiArgs<W> e = make_epi<W>(w, false, mask, f, accum);
                launch_vec_writeback<W>(e, (const W *)tc, tpres, n_out);
            });
            dfree(tc);
        }
        vec_set_unknown_nvals(w);
    };

    if (push) {
        const Csr *Mp = nullptr;
        int64_t *deg = nullptr, *offs = nullptr;
        int64_t nedges = 0;
        if (nu > 0) {
            vec_build_list(u);
            Mp = Aoverride ? Aoverride : get_csr(!use_transpose);
            deg = dalloc<int64_t>(nu + 1); offs = dalloc<int64_t>(nu + 1);
            GRB_LAUNCH("frontier_degrees", k_frontier_degrees, nblk(nu + 1), 256, 0, u->list, nu, Mp->rowptr, deg);
            size_t tb = 0;
            CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tb, deg, offs, nu + 1, g.stream));
            void *tmp = dmalloc(tb);
            CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, deg, offs, nu + 1, g.stream));
            g.launches += 2;
            dfree(tmp);
            nedges = read_i64(offs + nu);
            // direction by edges, not by vertices: a small frontier of hub vertices owns most of the matrix
            // (BFS level 2 on R-MAT: 4% of the vertices, 70% of the edges); the pull then reads less
            if (g.force_direction == 0 && !Aoverride && nedges * g.push_edge_ratio > Mp->nnz) push = false;
        }
        if (push) {
            t = dalloc<T>(n_out); tp = dalloc<uint8_t>(n_out);
            T ident = monoid_identity<T>(mon);
            GRB_LAUNCH("fill_identity", k_fill_identity<T>, ew_grid(n_out), 256, 0, t, tp, n_out, ident);
            if (nu > 0) {
                PushArgs<T> pa;
                pa.rowptr = Mp->rowptr; pa.col = Mp->colidx; pa.av = vals_of(const_cast<Csr *>(Mp)); pa.a_iso = Mp->iso;
                pa.u = (const T *)u_vals; pa.list = u->list; pa.nf = nu; pa.offs = offs; pa.nedges = nedges;
                pa.t = t; pa.tp = tp;
                pa.mask_p = (have_mask && f.structure) ? mask->present : nullptr; pa.comp = f.comp;
                if (!spmv_launch_push(mon, mul, tcode, &pa)) grb_throw(GrB_NOT_IMPLEMENTED, "semiring has no sm_100a kernel");
            }
            dfree(deg); dfree(offs);
            writeback(t, tp);
            dfree(t); dfree(tp);
            return;
        }
        dfree(deg); dfree(offs);
    }

    // ---- pull ----
    // (Aoverride given for push orientation only when the caller could not provide both; treat as pull source)
    Csr *M = Aoverride ? const_cast<Csr *>(Aoverride) : get_csr(use_transpose);
    const bool pull_any = any_mon && have_mask;
    PullArgs<T> a;
    a.rowptr = M->rowptr; a.col = M->colidx; a.av = vals_of(M); a.a_iso = M->iso;
    a.u = (const T *)u_vals; a.up = (nu == n_in) ? nullptr : u->present;
    a.nrows = M->nrows; a.nnz = M->nnz;
    a.carry_head = a.carry_own = nullptr; a.carry_head_has = a.carry_own_has = nullptr;
    a.tile_row = nullptr; a.ntiles = 0;
    a.hub_v = nullptr; a.hub_p = nullptr; a.hub_n = 0;
    // flagged-stream fast path: u full, and the matrix is pulled from repeatedly (iterative algorithms) so that
    // building its stream form (hub ranks, row-start flags) pays off
    // (the atomic-segment form also takes a sparse u: Bellman-Ford's `w(min) << w.vxm(A, min_plus)` before w fills up)
    const bool red_shape = !have_mask && !f.comp && (accum < 0 || accum == mon);
    const bool red_any = red_shape || g.hub_mode != 3;      // hub_mode 3: keep k_spmv_stream for non-monoid write-backs (A/B)
    bool use_stream = !pull_any && (a.up == nullptr || red_any) && g.hub_mode && M->nnz > 0 && M->ncols <= (int64_t)COL_MASK;
    // column panels for the atomic-segment form: when the vector is too large to stay in L2 (a multi-GPU row block
    // sees all n columns), the rows are split by column ranges and pulled panel by panel into the same output
    int npanels = 1;
    if (use_stream && red_any && g.panel_kb > 0 && M->nnz >= (1 << 20)) {
        const int64_t bytes = M->ncols * (int64_t)sizeof(T) * (a.up ? 2 : 1);
        npanels = (int)std::min<int64_t>(8, cdiv(bytes, g.panel_kb * 1024));
        npanels = (int)std::max<int64_t>(1, std::min<int64_t>(npanels, M->nnz >> 24));    // >= 16 M entries per panel: each
                                                                                          // panel costs a hub gather + a launch
    }
    if (use_stream) {
        M->pull_uses++;
        const bool build = g.hub_mode >= 2 || (M->pull_uses >= 2 && M->nnz >= (1 << 20)) || prog.counters != nullptr;
        if (npanels > 1) {
            if (build) {
                csr_ensure_panels(*M, npanels, uses_x ? (const void *)a.av : nullptr, tcode);
                for (int p = 0; p < M->n_panels; p++) csr_ensure_hubs(M->panels[p], hub_capacity<T>());
            }
            use_stream = M->n_panels == npanels && M->panel_code == (uses_x ? tcode : 100 + tcode);
            for (int p = 0; use_stream && p < M->n_panels; p++)
                use_stream = M->panels[p].nnz == 0 || (M->panels[p].hub_col != nullptr && M->panels[p].hub_n <= hub_capacity<T>());
        } else {
            if (build) csr_ensure_hubs(*M, hub_capacity<T>());
            use_stream = M->hub_col != nullptr && M->hub_n <= hub_capacity<T>();
        }
    }
    // ---- atomic-segment pull.  When the write-back is the monoid itself (or a plain overwrite) and there is no mask,
    // partial sums go straight into w; any other write-back (mask, replace, foreign accumulator, type cast) first
    // reduces into a temporary t and then runs the dense C-API write-back over it: 190 + 20 us against 380-450 us
    // for the scan-based k_spmv_stream + fix-up passes on the PageRank graph (tools/debug/masked_variants.py).
    if (prog.counters && !(use_stream && red_shape && fused && npanels == 1 && a.up == nullptr && nu > 0))
        grb_throw(GrB_NOT_IMPLEMENTED, "progress reporting (GxB_b200_next_product_progress) needs the unmasked pull `w(monoid) << A.mxv(u)` "
                                       "with u full, w of the semiring's type and a matrix that fits one column panel");
    if (use_stream && (red_shape || g.hub_mode != 3)) {
        const bool direct = fused && red_shape;
        const bool pingpong = direct && ((w == u) || (u_vals == w->vals));
        RedArgs<T> ra;
        ra.u = (const T *)u_vals; ra.up = a.up; ra.flags = (int)g.red_flags;
        ra.upairs = nullptr;
        ra.progress = nullptr; ra.tile_blocks = nullptr; ra.stamps = nullptr;
        if (prog.counters) {
            // warps of the launch below (grb_spmv_inst.inc: launch_red_v): min(ceil(tiles / W), SMs) CTAs of W warps
            const int64_t W = sizeof(T) > 4 ? RED_WARPS_8B : 32, want = cdiv(M->n_stiles, W);
            csr_ensure_tile_blocks(*M, prog.rpb, prog.nblocks, (want < g.sm_count ? (want > 0 ? want : 1) : g.sm_count) * W);
            ra.progress = prog.counters; ra.tile_blocks = M->tile_blocks;
            ra.stamps = prog.counters + prog.nblocks;       // the caller's counter array has two spare words behind the blocks
            if (prog.expected) d2d(prog.expected, M->tile_expected, (size_t)prog.nblocks * 8);
        }
        if (ra.up) {      // sparse u: interleave (value, presence) so that one gather fetches both
            pair_bits_t<T> *pairs = (pair_bits_t<T> *)dmalloc((size_t)n_in * sizeof(pair_bits_t<T>));
            GRB_LAUNCH("pack_pairs", k_pack_pairs<T>, ew_grid(n_in), 256, 0, ra.u, ra.up, n_in, pairs);
            ra.upairs = pairs;
        }
        const T ident = monoid_identity<T>(mon);
        const bool overwrite = !direct || accum < 0;
        // (a vector wrapped over caller-owned memory is full by contract, GxB_b200_Vector_wrap_device: its count is never lost)
        const bool w_full = direct && (w->external_vals || (w->nvals_known && w->nvals == n_out));
        if (direct) {
            if (pingpong) {
                vec_ensure_alt(w);
                if (!overwrite) { d2d(w->vals_alt, w->vals, (size_t)n_out * sizeof(T)); d2d(w->present_alt, w->present, (size_t)n_out); }
                ra.w = (T *)w->vals_alt; ra.wp = w->present_alt;
            } else { ra.w = (T *)w->vals; ra.wp = w->present; }
        } else {
            t = dalloc<T>(n_out); tp = dalloc<uint8_t>(n_out);
            ra.w = t; ra.wp = tp;
        }
        if (overwrite || !w_full)
            GRB_LAUNCH("red_prefill", k_red_prefill<T>, ew_grid(n_out), 256, 0, ra.w, ra.wp, n_out, ident, (int)overwrite);
        if (!overwrite && w_full) ra.wp = nullptr;     // stays full: no presence traffic at all
        bool ok = true;
        const int nparts = npanels > 1 ? M->n_panels : 1;
        for (int p = 0; ok && p < nparts; p++) {
            Csr *Mp = npanels > 1 ? &M->panels[p] : M;
            if (Mp->nnz == 0) continue;
            ra.pcol = Mp->hub_col; ra.av = npanels > 1 ? (const T *)Mp->vals : a.av; ra.a_iso = Mp->iso;
            ra.hub_n = Mp->hub_n; ra.nzrow = Mp->nzrow; ra.chunk_first = Mp->chunk_first; ra.nnz = Mp->nnz; ra.ntiles = Mp->n_stiles;
            // the hub values of this call live in a buffer kept with the matrix: no stream-ordered free between the product and
            // whatever the caller launches right behind it (the fused peer exchange relies on being the next node in the stream)
            if (Mp->hub_buf_bytes < (int64_t)ra.hub_n * (int64_t)sizeof(T) || !Mp->hub_vbuf) {
                dfree(Mp->hub_vbuf); dfree(Mp->hub_pbuf);
                Mp->hub_buf_bytes = (int64_t)ra.hub_n * 8;
                Mp->hub_vbuf = dmalloc((size_t)Mp->hub_buf_bytes); Mp->hub_pbuf = dalloc<uint8_t>(ra.hub_n);
            }
            T *hv = (T *)Mp->hub_vbuf; uint8_t *hp = Mp->hub_pbuf;
            GRB_LAUNCH("gather_hubs", k_gather_hubs<T>, nblk(ra.hub_n), 256, 0, ra.u, ra.up, Mp->hub_ids, ra.hub_n, hv, hp);
            ra.hub_v = hv; ra.hub_p = hp;
            ok = spmv_launch_red(mon, mul, tcode, &ra);
        }
        dfree((void *)ra.upairs);
        if (!ok) { dfree(t); dfree(tp); grb_throw(GrB_NOT_IMPLEMENTED, "semiring has no sm_100a kernel"); }
        if (direct) {
            if (pingpong) swap_pingpong(w);
            if (overwrite && !ra.up && npanels == 1) { w->nvals = M->n_nzrow; w->nvals_known = true; w->list_valid = false; }
            else if (!overwrite && w_full) { w->list_valid = false; }
            else { count_present(w->present, n_out, w->d_nvals); vec_set_unknown_nvals(w); }
        } else {
            writeback(t, tp);
            dfree(t); dfree(tp);
        }
        return;
    }
    StreamArgs<T> sa;
    if (use_stream) {
        sa.rowptr = M->rowptr; sa.pcol = M->hub_col; sa.av = a.av; sa.a_iso = M->iso;
        sa.u = (const T *)u_vals; sa.hub_n = M->hub_n;
        sa.nzrow = M->nzrow; sa.tile_first = M->tile_first; sa.nrows = M->nrows; sa.nnz = M->nnz; sa.ntiles = M->n_stiles;
        T *hv = dalloc<T>(sa.hub_n); uint8_t *hp = dalloc<uint8_t>(sa.hub_n);
        GRB_LAUNCH("gather_hubs", k_gather_hubs<T>, nblk(sa.hub_n), 256, 0, sa.u, (const uint8_t *)nullptr, M->hub_ids, sa.hub_n, hv, hp);
        sa.hub_v = hv; a.hub_v = hv; a.hub_p = hp;
        sa.carry_head = dalloc<T>(2 * sa.ntiles); sa.carry_own = sa.carry_head + sa.ntiles;
        a.carry_head = sa.carry_head;
        sa.open_list = dalloc<int32_t>(sa.ntiles + 1); sa.open_count = (int *)(sa.open_list + sa.ntiles);
        dzero(sa.open_count, sizeof(int));
    }
    if (!pull_any && !use_stream) {
        csr_ensure_tiles(*M, SPMV_TILE);
        a.tile_row = M->tile_row; a.ntiles = M->ntiles;
        a.carry_head = dalloc<T>(2 * M->ntiles); a.carry_own = a.carry_head + M->ntiles;
        a.carry_head_has = dalloc<uint8_t>(2 * M->ntiles); a.carry_own_has = a.carry_head_has + M->ntiles;
    }
    auto launch = [&](const EpiArgs<T> &e) {
        bool ok;
        if (use_stream) {
            sa.epi = e;
            ok = spmv_launch_stream(mon, mul, tcode, &sa);
            if (ok && M->n_nzrow < M->nrows)
                GRB_LAUNCH("stream_empty_rows", k_stream_empty_rows<T>, ew_grid(M->nrows), 256, 0, M->rowptr, M->nrows, e);
        } else {
            a.epi = e;
            ok = spmv_launch_pull(mon, mul, tcode, &a, pull_any ? PULL_ANY : PULL_STREAM);
        }
        if (!ok) grb_throw(GrB_NOT_IMPLEMENTED, "semiring has no sm_100a kernel");
    };
    if (fused) {
        // w aliasing u (e.g. w(min) << w.vxm(A)): write to the ping-pong spare, then swap
        const bool pingpong = (w == u) || (u_vals == w->vals);
        dzero(w->d_nvals, sizeof(int64_t));
        launch(make_epi<T>(w, pingpong, mask, f, accum));
        if (pingpong) swap_pingpong(w);
        vec_set_unknown_nvals(w);
    } else {
        // compute t in the semiring's type, then cast + write back in w's type
        t = dalloc<T>(n_out); tp = dalloc<uint8_t>(n_out);
        dzero(tp, (size_t)n_out);
        EpiArgs<T> e;
        e.w_in = t; e.wp_in = tp; e.w_out = t; e.wp_out = tp;
        e.mask_p = nullptr; e.mask_v = nullptr; e.mask_vsize = 0; e.mask_fp = 0; e.comp = 0; e.replace = 0; e.accum = -1;
        e.count = (unsigned long long *)(g.d_scratch + 16);
        if (pull_any) { e.mask_p = mask->present; e.mask_v = mask->vals; e.mask_vsize = f.structure ? 0 : (int)mask->type->size;
                        e.mask_fp = (mask->type->code == TC_FP32 || mask->type->code == TC_FP64); e.comp = f.comp; }
        launch(e);
        writeback(t, tp);
        dfree(t); dfree(tp);
    }
    if (use_stream) dfree(sa.open_list);
    if (!pull_any) { dfree(a.carry_head); dfree(a.carry_head_has); dfree((void *)a.hub_v); dfree((void *)a.hub_p); }
}

// shared body of GrB_mxv / GrB_vxm
static void vector_product(GB_Object *w, GB_Object *mask, const GrB_BinaryOp accum, const GrB_Semiring sr,
                           GB_Object *A, GB_Object *u, const DescFlags &f, bool is_vxm) {
    require_init();
    ProgressReq prog;
    prog.counters = g.progress_dev; prog.expected = g.progress_expected_dev; prog.nblocks = g.progress_blocks; prog.rpb = g.progress_rpb;
    g.progress_dev = nullptr; g.progress_expected_dev = nullptr;       // consumed by this product, whatever happens to it
    if (!sr) grb_throw(GrB_NULL_POINTER, "semiring is NULL");
    if (w->kind == K_MATRIX || u->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "w and u must be vectors");
    if (mask && mask->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "mask must be a vector");
    const bool tran = is_vxm ? f.tran1 : f.tran0;
    // A may be a vector cast to an n x 1 matrix (Vector.inner, reference core/vector.py:1641-1648)
    GB_Object tmpA;
    bool A_is_tmp = false;
    if (A->kind != K_MATRIX) {
        tmpA.kind = K_MATRIX; tmpA.type = A->type; tmpA.nrows = A->nrows; tmpA.ncols = 1;
        tmpA.csr = vec_as_csr_column(A);
        A = &tmpA; A_is_tmp = true;
    }
    struct Cleanup { GB_Object *o; bool on; ~Cleanup() { if (on) { csr_free(o->csr); if (o->csc_valid) csr_free(o->csc); o->magic = 0; } } } cl{&tmpA, A_is_tmp};
    const int64_t ar = tran ? A->ncols : A->nrows, ac = tran ? A->nrows : A->ncols;   // effective shape
    const int64_t n_in = is_vxm ? ar : ac, n_out = is_vxm ? ac : ar;
    if (u->nrows != n_in || w->nrows != n_out || (mask && mask->nrows != n_out)) {
        char b[256];
        snprintf(b, sizeof b, "dimension mismatch: %s with A %lldx%lld%s, u size %lld, w size %lld%s",
                 is_vxm ? "vxm" : "mxv", (long long)A->nrows, (long long)A->ncols, tran ? " (transposed)" : "",
                 (long long)u->nrows, (long long)w->nrows, mask ? ", mask" : "");
        grb_throw(GrB_DIMENSION_MISMATCH, b);
    }
    int acc = -1;
    if (accum) {
        op_need_kernel(accum->op, accum->name);
        if (op_is_cmp(accum->op)) grb_throw(GrB_DOMAIN_MISMATCH, "comparison operators cannot accumulate into w");
        acc = accum->op;
    }
    op_need_kernel(sr->mon, sr->name);
    int tcode = sr->tcode;
    const int mon = sr->mon;
    int mul = sr->mul;
    // the kernels multiply (matrix entry, vector entry); vxm's multiply is (vector entry, matrix entry): swap the
    // operand-selecting and the non-commutative multiplies
    if (is_vxm) mul = (mul == OP_FIRST) ? OP_SECOND : (mul == OP_SECOND) ? OP_FIRST : (mul == OP_DIV) ? OP_RDIV : mul;
    // INT8/INT16/UINT8/UINT16 semirings run in the 32-bit type of the same signedness (grb_internal.cuh)
    int narrow = -1;
    if (tc_is_narrow(tcode)) { narrow = tcode; mul = narrow_mul(mon, mul, narrow); tcode = tc_wide_of(narrow); }
    // pull reads rows of (is_vxm != tran) ? A' : A
    const bool use_transpose = (is_vxm != tran);
    // reconcile operand types with the semiring's type (cast copies; the common case needs none)
    void *u_cast = nullptr;
    const void *u_vals = u->vals;
    if (u->type->code != tcode || narrow >= 0) {
        u_cast = dmalloc((size_t)(u->nrows > 0 ? u->nrows : 1) * kTypeSize[tcode]);
        if (narrow >= 0 && u->type->code != narrow) {
            void *un = dmalloc((size_t)(u->nrows > 0 ? u->nrows : 1) * kTypeSize[narrow]);
            cast_array(u->vals, u->type->code, un, narrow, u->nrows);
            cast_array(un, narrow, u_cast, tcode, u->nrows);
            dfree(un);
        } else cast_array(u->vals, u->type->code, u_cast, tcode, u->nrows);
        u_vals = u_cast;
    }
    // (A's values are cast inside product_typed, once per matrix: csr_vals_as)
    try {
        GRB_TYPE_SWITCH(tcode, T, product_typed<T>(w, mask, acc, mon, mul, tcode, A, use_transpose, nullptr, u, u_vals, f, narrow, prog));
    } catch (...) { dfree(u_cast); throw; }
    dfree(u_cast);
    if (g.blocking) stream_sync();
}

extern "C" {

GrB_Info GrB_mxv(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Semiring op,
                 const GrB_Matrix A, const GrB_Vector u, const GrB_Descriptor desc) {
    API_BEGIN
    GB_Object *W = obj_check(w, "w"), *Aobj = obj_check_lazy(A, "A"), *U = obj_check(u, "u");
    GB_Object *M = mask ? obj_check(mask, "mask") : nullptr;
    vector_product(W, M, accum, op, Aobj, U, desc_flags(desc), false);
    API_END(w)
}

GrB_Info GrB_vxm(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Semiring op,
                 const GrB_Vector u, const GrB_Matrix A, const GrB_Descriptor desc) {
    API_BEGIN
    GB_Object *W = obj_check(w, "w"), *Aobj = obj_check_lazy(A, "A"), *U = obj_check(u, "u");
    GB_Object *M = mask ? obj_check(mask, "mask") : nullptr;
    vector_product(W, M, accum, op, Aobj, U, desc_flags(desc), true);
    API_END(w)
}

// w<m> = accum(w, (+)_j A(i,j)): the row-wise reduce of the reference's aggregators (graphblas/core/operator/agg.py)
// is the product A (+).first 1 -- the multiply ignores the vector, so nothing is gathered and the pull kernels
// run as a pure stream.  BOOL monoids use their GrB semiring with the multiply's neutral element instead.
GrB_Info GrB_Matrix_reduce_Monoid(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Monoid monoid,
                                  const GrB_Matrix A, const GrB_Descriptor desc) {
    GrB_Vector ones = nullptr;
    API_BEGIN
    require_init();
    if (!monoid) grb_throw(GrB_NULL_POINTER, "monoid is NULL");
    GB_Object *W = obj_check(w, "w"), *Aobj = obj_check_lazy(A, "A");
    GB_Object *M = mask ? obj_check(mask, "mask") : nullptr;
    if (Aobj->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "A must be a matrix");
    const DescFlags f = desc_flags(desc);
    op_need_kernel(monoid->op, monoid->name);
    const int tcode = monoid->tcode;
    GB_Semiring_opaque sr = {monoid->op, OP_FIRST, tcode, "reduce_rowwise"};
    bool fill = true;
    if (tcode == TC_BOOL) {
        switch (monoid->op) {
            case OP_LOR: sr.mul = OP_LAND; fill = true; break;
            case OP_LXOR: sr.mul = OP_LAND; fill = true; break;
            case OP_LAND: sr.mul = OP_LOR; fill = false; break;
            case OP_LXNOR: sr.mul = OP_LOR; fill = false; break;
            default: break;      // ANY_FIRST exists
        }
    }
    const GrB_Index n = f.tran0 ? Aobj->nrows : Aobj->ncols;
    static const GrB_Type types[TC_COUNT] = {GrB_BOOL, GrB_INT8, GrB_INT16, GrB_INT32, GrB_INT64, GrB_UINT8,
                                             GrB_UINT16, GrB_UINT32, GrB_UINT64, GrB_FP32, GrB_FP64};
    GrB_Info info = GrB_Vector_new(&ones, types[tcode], n);
    if (info == GrB_SUCCESS) info = GrB_Vector_assign_BOOL(ones, nullptr, nullptr, fill, GrB_ALL, n, nullptr);
    if (info != GrB_SUCCESS) grb_throw(info, "reduce: cannot build the operand vector");
    try {
        vector_product(W, M, accum, &sr, Aobj, (GB_Object *)ones, f, false);
    } catch (...) { GrB_Vector_free(&ones); throw; }
    GrB_Vector_free(&ones);
    API_END(w)
}

}  // extern "C"
