// gather_host.inl — host side of b200j_gather (included inside b200join.cu's anonymous namespace).
// take_multiple_record_batch (src/shared/shared.rs:83-92): build columns (keys, then payload) at build_idx, then the probe
// batch's columns (keys, then payload) at probe_idx. LeftSemi/LeftAnti output only the build side, RightSemi/RightAnti
// only the probe side (build_join_schema, probe_lookup_implementation.rs:77-94).

struct GatheredOwner {
    std::vector<b200j_out_column> cols;
    std::vector<void *> bufs;          // every buffer handed out (device: big_free; host: pinned_give)
    int32_t mem = B200J_MEM_DEVICE;
};

struct GatherCtx {
    H *h;
    int64_t n;
    const uint64_t *bidx;              // device
    const uint32_t *pidx;              // device
    std::vector<void *> temps;
    GatheredOwner *own;
    unsigned long long *d_nulls;       // device counter per column (array)
    unsigned long long *d_nulls_scratch;   // GF_MAX_COLS counters of the multi-column launch in flight
};

void *gather_dev_alloc(GatherCtx &g, size_t bytes) {
    void *p = nullptr;
    CU(big_alloc(g.h, &p, std::max<size_t>(bytes, 256)));
    g.own->bufs.push_back(p);
    return p;
}

template <typename IdxT>
void gather_one_column(GatherCtx &g, int col, int type, const GatherSrc &src, const IdxT *idx) {
    H *h = g.h;
    const int64_t n = g.n;
    b200j_out_column &oc = g.own->cols[col];
    oc = b200j_out_column{};
    const size_t vwords = (size_t)((n + 31) / 32 + 1);
    uint32_t *dvalid = (uint32_t *)gather_dev_alloc(g, vwords * 4);
    oc.validity = (uint8_t *)dvalid;
    unsigned long long *nulls = g.d_nulls + col;
    const int grid = grid_for(h, n, 256, 8);
    if (type != B200J_T_UTF8) {
        void *dst = gather_dev_alloc(g, (size_t)n * type + 16);
        oc.values = dst;
        oc.values_bytes = n * type;
        if (n > 0) {
            switch (type) {
                case 1: gather_fixed_kernel<uint8_t, IdxT><<<grid, 256, 0, h->stream>>>(src, idx, n, (uint8_t *)dst, dvalid, nulls); break;
                case 2: gather_fixed_kernel<uint16_t, IdxT><<<grid, 256, 0, h->stream>>>(src, idx, n, (uint16_t *)dst, dvalid, nulls); break;
                case 4: gather_fixed_kernel<uint32_t, IdxT><<<grid, 256, 0, h->stream>>>(src, idx, n, (uint32_t *)dst, dvalid, nulls); break;
                case 8: gather_fixed_kernel<unsigned long long, IdxT><<<grid, 256, 0, h->stream>>>(src, idx, n, (unsigned long long *)dst, dvalid, nulls); break;
                default: gather_fixed_kernel<ulonglong2, IdxT><<<grid, 256, 0, h->stream>>>(src, idx, n, (ulonglong2 *)dst, dvalid, nulls); break;
            }
            h->st.kernel_launches++;
            CU(cudaGetLastError());
        }
        return;
    }
    // Utf8: lengths -> exclusive scan -> offsets, then the byte copy
    int32_t *ooff = (int32_t *)gather_dev_alloc(g, (size_t)(n + 1) * 4 + 16);
    oc.offsets = ooff;
    if (n == 0) {
        CU(cudaMemsetAsync(ooff, 0, 4, h->stream));
        oc.values = gather_dev_alloc(g, 256);
        return;
    }
    int64_t *len = nullptr, *start = nullptr, *bsum = nullptr, *dtotal = nullptr;
    const int64_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    CU(big_alloc(h, &len, (size_t)n * 8)); g.temps.push_back(len);
    CU(big_alloc(h, &start, (size_t)n * 8)); g.temps.push_back(start);
    CU(big_alloc(h, &bsum, (size_t)(n_tiles + 1) * 8)); g.temps.push_back(bsum);
    dtotal = bsum + n_tiles;
    gather_utf8_len_kernel<IdxT><<<grid, 256, 0, h->stream>>>(src, idx, n, len, dvalid, nulls);
    scan_block_sums_kernel<<<(unsigned)n_tiles, SCAN_THREADS, 0, h->stream>>>(len, n, bsum);
    scan_serial_kernel<<<1, 1024, 0, h->stream>>>(bsum, n_tiles, dtotal);
    scan_apply_kernel<<<(unsigned)n_tiles, SCAN_THREADS, 0, h->stream>>>(len, n, bsum, start, ooff);
    h->st.kernel_launches += 4;
    CU(cudaGetLastError());
    int64_t total = 0;
    CU(cudaMemcpyAsync(&total, dtotal, 8, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    {   // out32[n] = total (the apply kernel only covers it when n is not a multiple of the scan tile)
        const int32_t t32 = (int32_t)std::min<int64_t>(total, 0x7fffffffLL);
        CU(cudaMemcpyAsync(ooff + n, &t32, 4, cudaMemcpyHostToDevice, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    REQUIRE(total <= 0x7fffffffLL, B200J_ERR_UNSUPPORTED, "gathered Utf8 column exceeds 2 GiB (32-bit Arrow offsets): gather fewer rows per call");
    char *dst = (char *)gather_dev_alloc(g, (size_t)total + 16);
    oc.values = dst;
    oc.values_bytes = total;
    if (total > 0) {
        double avg = (double)total / (double)n;                // bytes per gathered value (NULL rows count as 0)
        if (h->utf8_lanes > 0) avg = h->utf8_lanes <= 4 ? 1.0 : h->utf8_lanes <= 8 ? 10.0 : h->utf8_lanes <= 16 ? 40.0 : 100.0;      // B200J_UTF8_LANES: tuning
        if (avg <= 6.0) gather_utf8_copy_kernel<IdxT, 4><<<grid_for(h, n * 4, 256, 16), 256, 0, h->stream>>>(src, idx, n, start, len, dst);
        else if (avg <= 16.0) gather_utf8_copy_kernel<IdxT, 8><<<grid_for(h, n * 8, 256, 16), 256, 0, h->stream>>>(src, idx, n, start, len, dst);
        else if (avg <= 48.0) gather_utf8_copy_kernel<IdxT, 16><<<grid_for(h, n * 16, 256, 16), 256, 0, h->stream>>>(src, idx, n, start, len, dst);
        else gather_utf8_copy_kernel<IdxT, 32><<<grid_for(h, n * 32, 256, 16), 256, 0, h->stream>>>(src, idx, n, start, len, dst);
        h->st.kernel_launches++;
        CU(cudaGetLastError());
    }
}

// output buffers of one fixed-width column: values + validity words (a few spare words: the kernels write whole words)
void gather_out_fixed(GatherCtx &g, int col, int type, void **dst, uint32_t **dvalid) {
    b200j_out_column &oc = g.own->cols[col];
    oc = b200j_out_column{};
    const size_t vwords = (size_t)((g.n + 31) / 32 + 4);
    *dvalid = (uint32_t *)gather_dev_alloc(g, vwords * 4);
    oc.validity = (uint8_t *)*dvalid;
    *dst = gather_dev_alloc(g, (size_t)g.n * type + 16);
    oc.values = *dst;
    oc.values_bytes = g.n * type;
}

// All columns of one side: the fixed-width ones go through gather_fixed_multi_kernel, up to GF_MAX_COLS per launch (the
// index array is read once per launch, not once per column); Utf8 columns take the three-pass path above.
struct SideCol {
    int type;
    GatherSrc src;
};
constexpr int SIDECOL_SKIP = -12345;

// Slot-ordered copies of the fixed-width build payload columns (B200J_OUT_GATHER_HINT, see gather.cuh): built once per build.
// Returns false when they cannot be had (duplicate build keys, slot indices beyond 32 bits, nothing to copy).
bool ensure_slot_cols(H *h) {
    const int nbp = h->cfg.n_build_payload;
    if (h->slot_cols_ready) return true;
    if (h->layout != B200J_LAYOUT_RADIX || h->st.duplicate_rows != 0 || h->n_build == 0 || h->n_slots + 2 >= 0xffffffffULL) return false;
    bool any = false;
    for (int i = 0; i < nbp; i++) any |= h->bpayload[i].type != B200J_T_UTF8;
    if (!any) return false;
    h->slot_cols.resize(nbp);
    const uint64_t total = h->n_slots + 2;
    for (int i = 0; i < nbp; i++) {                      // buffers first (kept across builds): the timer below covers the kernels only
        const ColStore &cs = h->bpayload[i];
        if (cs.type == B200J_T_UTF8) continue;
        dev_alloc(h, h->slot_cols[i].values, (size_t)total * cs.type + 32);
        if (cs.has_nulls) dev_alloc(h, h->slot_cols[i].validity, (size_t)((total + 31) / 32 + 1) * 4);
    }
    PhaseTimer tm(h, &h->st.ms_slot_order);
    SlotOrderCols batch{};
    auto flush = [&]() {
        if (batch.n == 0) return;
        slot_order_kernel<<<grid_for(h, (int64_t)total, 256, 16), 256, 0, h->stream>>>((const Slot *)h->slots.p, h->n_slots, (const Slot *)h->zero_slot.p, batch);
        h->st.kernel_launches++;
        CU(cudaGetLastError());
        batch = SlotOrderCols{};
    };
    for (int i = 0; i < nbp; i++) {
        const ColStore &cs = h->bpayload[i];
        if (cs.type == B200J_T_UTF8) continue;
        H::SlotCol &sc = h->slot_cols[i];
        sc.has_nulls = cs.has_nulls;
        SlotOrderCol &c = batch.c[batch.n++];
        c.src = cs.values.p;
        c.src_valid = cs.has_nulls ? (const uint8_t *)cs.validity.p : nullptr;
        c.dst = sc.values.p;
        c.dst_valid = cs.has_nulls ? (uint32_t *)sc.validity.p : nullptr;
        c.width = cs.type;
        if (batch.n == GF_MAX_COLS) flush();
    }
    flush();
    tm.stop();
    h->slot_cols_ready = true;
    return true;
}
// skip_first: column first_col has been produced already (gather_keys_both_kernel). seg_ro: the pairs are a dense radix result
// that still carries its segment table -- the fixed-width columns are then gathered partition by partition.
template <typename IdxT>
void gather_side(GatherCtx &g, int first_col, const std::vector<SideCol> &sc, const IdxT *idx, const void *key_via_probe = nullptr,
                 bool skip_first = false, const ResultOwner *seg_ro = nullptr) {
    H *h = g.h;
    const int64_t n = g.n;
    GatherCols batch{};
    int batch_cols[GF_MAX_COLS];
    auto flush = [&]() {
        if (batch.n == 0) return;
        if (n > 0) {
            // the kernel's null counters are indexed by position in the batch: count into a scratch row, add up afterwards
            if (seg_ro != nullptr) {
                GatherSegArgs sg{};
                sg.seg = (const uint32_t *)seg_ro->seg;
                sg.block_base = sg.seg + (size_t)seg_ro->seg_blocks * (seg_ro->seg_P + 2);
                sg.n_blocks = seg_ro->seg_blocks;
                sg.P = seg_ro->seg_P;
                sg.n_units = seg_ro->seg_blocks * (seg_ro->seg_P + 1);
                sg.next = (unsigned int *)(g.d_nulls_scratch + GF_MAX_COLS);
                for (int k = 0; k < batch.n; k++)          // all-valid bitmaps: NULL rows clear their bits
                    CU(cudaMemsetAsync(batch.c[k].dst_valid, 0xff, (size_t)((n + 31) / 32) * 4, h->stream));
                const int64_t n_claims = ((int64_t)sg.n_units + GS_UNITS - 1) / GS_UNITS;
                const int grid = grid_for(h, n_claims, 8, 8);
                gather_seg_multi_kernel<IdxT><<<grid, 256, 0, h->stream>>>(batch, idx, n, sg, g.d_nulls_scratch);
                h->st.gather_seg_launches++;
            } else {
                const int grid = grid_for(h, (n + 127) / 128, 8, 8);
                gather_fixed_multi_kernel<IdxT><<<grid, 256, 0, h->stream>>>(batch, idx, n, g.d_nulls_scratch);
            }
            h->st.kernel_launches++;
            CU(cudaGetLastError());
            for (int k = 0; k < batch.n; k++)
                CU(cudaMemcpyAsync(g.d_nulls + batch_cols[k], g.d_nulls_scratch + k, 8, cudaMemcpyDeviceToDevice, h->stream));
            CU(cudaMemsetAsync(g.d_nulls_scratch, 0, (GF_MAX_COLS + 1) * 8, h->stream));
        }
        batch = GatherCols{};
    };
    for (size_t k = skip_first ? 1 : 0; k < sc.size(); k++) {
        const int col = first_col + (int)k;
        if (sc[k].type == SIDECOL_SKIP) continue;            // produced by another launch
        if (sc[k].type == B200J_T_UTF8) {
            gather_one_column<IdxT>(g, col, sc[k].type, sc[k].src, idx);
            continue;
        }
        b200j_out_column &oc = g.own->cols[col];
        oc = b200j_out_column{};
        const size_t vwords = (size_t)((n + 31) / 32 + 4);
        uint32_t *dvalid = (uint32_t *)gather_dev_alloc(g, vwords * 4);
        oc.validity = (uint8_t *)dvalid;
        void *dst = gather_dev_alloc(g, (size_t)n * sc[k].type + 16);
        oc.values = dst;
        oc.values_bytes = n * sc[k].type;
        if (k == 0 && key_via_probe != nullptr && n > 0) {
            // the build side's single fixed-width key column: taken from the probe batch's key column (gather_key_via_probe_kernel)
            const int grid = grid_for(h, (n + 127) / 128, 8, 8);
            if (sc[k].type == 8)
                gather_key_via_probe_kernel<unsigned long long, IdxT><<<grid, 256, 0, h->stream>>>(
                    idx, g.pidx, n, (const unsigned long long *)sc[k].src.values, sc[k].src.validity, (const unsigned long long *)key_via_probe,
                    (unsigned long long *)dst, dvalid, g.d_nulls + col);
            else
                gather_key_via_probe_kernel<uint32_t, IdxT><<<grid, 256, 0, h->stream>>>(
                    idx, g.pidx, n, (const uint32_t *)sc[k].src.values, sc[k].src.validity, (const uint32_t *)key_via_probe, (uint32_t *)dst, dvalid,
                    g.d_nulls + col);
            h->st.kernel_launches++;
            CU(cudaGetLastError());
            continue;
        }
        GatherCol &gc = batch.c[batch.n];
        gc.values = sc[k].src.values;
        gc.validity = sc[k].src.validity;
        gc.bit_offset = sc[k].src.bit_offset;
        gc.dst = dst;
        gc.dst_valid = dvalid;
        gc.width = sc[k].type;
        batch_cols[batch.n] = col;
        if (++batch.n == GF_MAX_COLS) flush();
    }
    flush();
}

// stages one column of a (host or device) batch; returns the device-side source descriptor
GatherSrc gather_stage_column(GatherCtx &g, const b200j_column &c, int type, int64_t n_rows, int mem) {
    H *h = g.h;
    GatherSrc s{};
    if (n_rows == 0) return s;
    if (mem == B200J_MEM_DEVICE) {
        s.values = c.values;
        s.offsets = c.offsets;
        s.validity = c.validity;
        s.bit_offset = c.bit_offset;
        return s;
    }
    if (type != B200J_T_UTF8) {
        void *d = nullptr;
        CU(big_alloc(h, &d, (size_t)n_rows * type + 16)); g.temps.push_back(d);
        CU(cudaMemcpyAsync(d, c.values, (size_t)n_rows * type, cudaMemcpyHostToDevice, h->stream));
        s.values = d;
    } else {
        REQUIRE(c.offsets != nullptr, B200J_ERR_INVALID, "Utf8 column without offsets buffer");
        const int32_t o0 = c.offsets[0], o1 = c.offsets[n_rows];
        void *doff = nullptr, *dd = nullptr;
        CU(big_alloc(h, &doff, (size_t)(n_rows + 1) * 4)); g.temps.push_back(doff);
        CU(cudaMemcpyAsync(doff, c.offsets, (size_t)(n_rows + 1) * 4, cudaMemcpyHostToDevice, h->stream));
        CU(big_alloc(h, &dd, (size_t)std::max<int64_t>(o1 - o0, 1) + 16)); g.temps.push_back(dd);
        if (o1 > o0) CU(cudaMemcpyAsync(dd, (const char *)c.values + o0, (size_t)(o1 - o0), cudaMemcpyHostToDevice, h->stream));
        s.offsets = doff;
        s.values = (const char *)dd - o0;          // offsets stay absolute
    }
    int64_t boff = 0;
    s.validity = stage_validity(h, c, n_rows, B200J_MEM_HOST, g.temps, &boff);
    s.bit_offset = boff;
    return s;
}

// staged_probe: the probe side comes from a partition's coalescing stage (b200j_probe_push) instead of a caller's batch
void do_gather_impl(H *h, const b200j_result *pairs, const b200j_batch *probe_batch, uint32_t want, b200j_gathered *out,
                    const std::vector<SideCol> *staged_probe) {
    REQUIRE(h->built, B200J_ERR_INVALID, "gather before the build side finished");
    REQUIRE(pairs != nullptr && out != nullptr && out->struct_size >= sizeof(b200j_gathered), B200J_ERR_INVALID, "gather: missing argument");
    const int jt = h->cfg.join_type;
    const bool want_build = !(jt == B200J_RIGHT_SEMI || jt == B200J_RIGHT_ANTI);
    const bool want_probe = !(jt == B200J_LEFT_SEMI || jt == B200J_LEFT_ANTI);
    const int nk = h->cfg.n_keys, nbp = h->cfg.n_build_payload, npp = h->cfg.n_probe_payload;
    if (probe_batch) {
        check_batch(h, probe_batch, npp);
        REQUIRE(probe_batch->n_payload == npp, B200J_ERR_INVALID, "gather needs every configured probe payload column");
    }
    const int64_t n = pairs->n;
    const bool pairs_probe_order = pairs->struct_size >= offsetof(b200j_result, probe_order) + sizeof(int32_t) && pairs->probe_order != 0;
    REQUIRE(!pairs_probe_order || (probe_batch == nullptr && staged_probe == nullptr), B200J_ERR_INVALID,
            "gather of a probe-order result: the probe batch's own columns are the output's probe columns (row i belongs to build_idx[i]); "
            "pass probe_batch = NULL to gather the build side");
    GatheredOwner *own = new GatheredOwner();
    GatherCtx g{h, n, nullptr, nullptr, {}, own, nullptr, nullptr};
    try {
        const int n_cols = (want_build ? nk + nbp : 0) + (want_probe ? nk + npp : 0);
        own->cols.resize(std::max(n_cols, 1));
        // index arrays on the device
        auto stage_idx = [&](const void *p, size_t w) -> const void * {
            if (!p || n == 0) return nullptr;
            if (pairs->mem == B200J_MEM_DEVICE) return p;
            void *d = nullptr;
            CU(big_alloc(h, &d, (size_t)n * w)); g.temps.push_back(d);
            CU(cudaMemcpyAsync(d, p, (size_t)n * w, cudaMemcpyHostToDevice, h->stream));
            return d;
        };
        const int bwidth = (pairs->struct_size >= offsetof(b200j_result, build_idx_width) + sizeof(int32_t) && pairs->build_idx_width == 4) ? 4 : 8;
        g.bidx = (const uint64_t *)stage_idx(pairs->build_idx, (size_t)bwidth);
        g.pidx = (const uint32_t *)stage_idx(pairs->probe_idx, 4);
        void *dn = nullptr;
        CU(big_alloc(h, &dn, (size_t)(n_cols + 2 + GF_MAX_COLS) * 8)); g.temps.push_back(dn);      // + the segment gather's claim counter
        CU(cudaMemsetAsync(dn, 0, (size_t)(n_cols + 2 + GF_MAX_COLS) * 8, h->stream));
        g.d_nulls = (unsigned long long *)dn;
        g.d_nulls_scratch = g.d_nulls + n_cols + 1;
        if (pairs->mem == B200J_MEM_DEVICE && want_build && n > 0) h->gather_seen = true;     // later device probes of this handle keep what the gather wants
        // a dense radix result that still carries its segment table: build columns partition by partition
        const ResultOwner *seg_ro = nullptr;
        if (want_build && h->gather_seg && pairs->mem == B200J_MEM_DEVICE && pairs->_owner != nullptr && h->seg_owners.count(pairs->_owner) != 0) {
            const ResultOwner *cand = (const ResultOwner *)pairs->_owner;
            if (cand->seg != nullptr && cand->seg_n == n && cand->build_idx == (const void *)pairs->build_idx && cand->bwidth == bwidth && n > 0)
                seg_ro = cand;
        }
        // B200J_OUT_GATHER_HINT: the slot-ordered payload copies are made once per build, outside the gather's own timing (ms_slot_order)
        const bool slot_ok = seg_ro != nullptr && seg_ro->aux_slot != nullptr && nbp > 0 && ensure_slot_cols(h);
        PhaseTimer tm(h, &h->st.ms_gather);
        int col = 0;
        bool keys_done = false;
        // the probe side's column descriptors first: the build side's key column is taken from the probe key column when it can be
        std::vector<SideCol> psc;
        const bool have_probe = staged_probe != nullptr || (probe_batch && probe_batch->n_rows > 0);
        if (want_probe) {
            for (int i = 0; i < nk + npp && staged_probe == nullptr; i++) {
                const int type = i < nk ? h->key_widths[i] : h->probe_payload_types[i - nk];
                GatherSrc s{};
                if (have_probe)
                    s = gather_stage_column(g, i < nk ? probe_batch->keys[i] : probe_batch->payload[i - nk], type, probe_batch->n_rows, probe_batch->mem);
                psc.push_back(SideCol{type, s});
            }
            if (staged_probe) psc = *staged_probe;
        }
        if (want_build) {
            std::vector<SideCol> sc;
            for (int i = 0; i < nk + nbp; i++) {
                const ColStore &cs = i < nk ? h->bkeys[i] : h->bpayload[i - nk];
                GatherSrc s{};
                s.values = cs.values.p;
                s.offsets = cs.offsets.p;
                s.offsets64 = 1;
                s.validity = cs.has_nulls ? (const uint8_t *)cs.validity.p : nullptr;
                if (h->n_build == 0) s.values = s.offsets = nullptr;
                sc.push_back(SideCol{cs.type, s});
            }
            // single fixed-width key, both sides in the output, a probe side to read from: build key := probe key of the pair
            const void *via = (h->key_via_probe && want_probe && have_probe && nk == 1 && !h->composite && g.pidx != nullptr && g.bidx != nullptr &&
                               (h->key_widths[0] == 4 || h->key_widths[0] == 8)) ? psc[0].src.values : nullptr;
            bool hinted = false;
            if (via != nullptr && n > 0 && seg_ro != nullptr && seg_ro->aux_keys != nullptr &&
                (seg_ro->aux_kw == h->key_widths[0] || (seg_ro->aux_kw == 4 && h->key_widths[0] == 8))) {
                // B200J_OUT_GATHER_HINT: both key columns are a sequential copy of the partitioned key stream
                const int kw = h->key_widths[0], pcol = nk + nbp;
                void *dst_b = nullptr, *dst_p = nullptr;
                uint32_t *val_b = nullptr, *val_p = nullptr;
                gather_out_fixed(g, col, kw, &dst_b, &val_b);
                gather_out_fixed(g, pcol, kw, &dst_p, &val_p);
                const int grid = grid_for(h, (n + 127) / 128, 8, 8);
                const void *nullsrc = seg_ro->aux_misses > 0 ? (const void *)g.bidx : nullptr;
                auto launch = [&](auto K_, auto T_, auto I_) {
                    using K = decltype(K_);
                    using T = decltype(T_);
                    using I = decltype(I_);
                    gather_keys_stream_kernel<K, T, I><<<grid, 256, 0, h->stream>>>((const K *)seg_ro->aux_keys, (const I *)nullsrc, n, (T *)dst_b, val_b,
                                                                                    (T *)dst_p, g.d_nulls + col);
                };
                if (kw == 8 && seg_ro->aux_kw == 4) {
                    if (bwidth == 4) launch((uint32_t)0, (unsigned long long)0, (uint32_t)0);
                    else launch((uint32_t)0, (unsigned long long)0, (uint64_t)0);
                } else if (kw == 8) {
                    if (bwidth == 4) launch((unsigned long long)0, (unsigned long long)0, (uint32_t)0);
                    else launch((unsigned long long)0, (unsigned long long)0, (uint64_t)0);
                } else {
                    if (bwidth == 4) launch((uint32_t)0, (uint32_t)0, (uint32_t)0);
                    else launch((uint32_t)0, (uint32_t)0, (uint64_t)0);
                }
                h->st.kernel_launches++;
                CU(cudaGetLastError());
                keys_done = true;
                hinted = true;
            } else if (via != nullptr && n > 0) {
                // ... and both key columns come out of ONE pass over the pairs (gather_keys_both_kernel)
                const int kw = h->key_widths[0], pcol = nk + nbp;
                void *dst_b = nullptr, *dst_p = nullptr;
                uint32_t *val_b = nullptr, *val_p = nullptr;
                gather_out_fixed(g, col, kw, &dst_b, &val_b);
                gather_out_fixed(g, pcol, kw, &dst_p, &val_p);
                const int grid = grid_for(h, (n + 127) / 128, 8, 8);
                auto launch = [&](auto T_, auto I_) {
                    using T = decltype(T_);
                    using I = decltype(I_);
                    gather_keys_both_kernel<T, I><<<grid, 256, 0, h->stream>>>(
                        (const I *)g.bidx, g.pidx, n, (const T *)sc[0].src.values, sc[0].src.validity, (const T *)psc[0].src.values, psc[0].src.validity,
                        psc[0].src.bit_offset, (T *)dst_b, val_b, (T *)dst_p, val_p, g.d_nulls + col, g.d_nulls + pcol);
                };
                if (kw == 8 && bwidth == 4) launch((unsigned long long)0, (uint32_t)0);
                else if (kw == 8) launch((unsigned long long)0, (uint64_t)0);
                else if (bwidth == 4) launch((uint32_t)0, (uint32_t)0);
                else launch((uint32_t)0, (uint64_t)0);
                h->st.kernel_launches++;
                CU(cudaGetLastError());
                keys_done = true;
            }
            // B200J_OUT_GATHER_HINT: fixed-width payload columns from their slot-ordered copies, indexed by the slot of every pair
            if (slot_ok) {
                std::vector<SideCol> ssc(sc.size(), SideCol{SIDECOL_SKIP, GatherSrc{}});
                for (int i = 0; i < nbp; i++) {
                    if (h->bpayload[i].type == B200J_T_UTF8) continue;
                    GatherSrc s{};
                    s.values = h->slot_cols[i].values.p;
                    s.validity = h->slot_cols[i].has_nulls ? (const uint8_t *)h->slot_cols[i].validity.p : nullptr;
                    ssc[nk + i] = SideCol{h->bpayload[i].type, s};
                    sc[nk + i].type = SIDECOL_SKIP;
                }
                gather_side<uint32_t>(g, col, ssc, (const uint32_t *)seg_ro->aux_slot, nullptr, false, seg_ro);
                hinted = true;
            }
            if (hinted) h->st.gather_hint_launches++;
            if (bwidth == 4) gather_side<uint32_t>(g, col, sc, reinterpret_cast<const uint32_t *>(g.bidx), keys_done ? nullptr : via, keys_done, seg_ro);     // B200J_OUT_BUILD_U32 pairs
            else gather_side<uint64_t>(g, col, sc, g.bidx, keys_done ? nullptr : via, keys_done, seg_ro);
            col += nk + nbp;
        }
        if (want_probe) {
            gather_side<uint32_t>(g, col, psc, have_probe ? g.pidx : nullptr, nullptr, keys_done);
            col += nk + npp;
        }
        tm.stop();
        // null counts; a column without nulls drops its validity buffer pointer (Arrow: NULL bitmap = no nulls)
        std::vector<unsigned long long> hn(n_cols + 1, 0);
        CU(cudaMemcpyAsync(hn.data(), dn, (size_t)n_cols * 8, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        for (int c = 0; c < n_cols; c++) {
            own->cols[c].null_count = (int64_t)hn[c];
            if (hn[c] == 0) own->cols[c].validity = nullptr;
        }
        if (want & B200J_OUT_HOST) {
            // move every buffer to pinned host memory
            std::vector<void *> host_bufs;
            for (int c = 0; c < n_cols; c++) {
                b200j_out_column &oc = own->cols[c];
                auto to_host = [&](void *d, size_t bytes) -> void * {
                    if (!d) return nullptr;
                    void *hp = pinned_take(h, std::max<size_t>(bytes, 1));
                    host_bufs.push_back(hp);
                    if (bytes) CU(cudaMemcpyAsync(hp, d, bytes, cudaMemcpyDeviceToHost, h->stream));
                    return hp;
                };
                oc.values = to_host(oc.values, (size_t)oc.values_bytes);
                oc.offsets = (int32_t *)to_host(oc.offsets, oc.offsets ? (size_t)(n + 1) * 4 : 0);
                oc.validity = (uint8_t *)to_host(oc.validity, (size_t)((n + 31) / 32) * 4);
            }
            CU(cudaStreamSynchronize(h->stream));
            for (void *p : own->bufs) big_free(h, p);
            own->bufs = host_bufs;
            own->mem = B200J_MEM_HOST;
        }
        free_temps(h, g.temps);
        out->mem = own->mem;
        out->n_rows = n;
        out->n_cols = n_cols;
        out->cols = own->cols.data();
        out->_owner = own;
    } catch (...) {
        free_temps(h, g.temps);
        for (void *p : own->bufs) big_free(h, p);
        delete own;
        throw;
    }
}

void do_gather(H *h, const b200j_result *pairs, const b200j_batch *probe_batch, uint32_t want, b200j_gathered *out) {
    do_gather_impl(h, pairs, probe_batch, want, out, nullptr);
}

void do_gathered_release(H *h, b200j_gathered *gth) {
    if (!gth || !gth->_owner) return;
    GatheredOwner *own = (GatheredOwner *)gth->_owner;
    for (void *p : own->bufs) {
        if (own->mem == B200J_MEM_DEVICE) big_free(h, p);
        else pinned_give(h, p);
    }
    delete own;
    gth->_owner = nullptr;
    gth->cols = nullptr;
    gth->n_cols = 0;
    gth->n_rows = 0;
}


// ---- coalescing feeder ------------------------------------------------------------------------------------------------
H::ProbeStage &stage_of(H *h, int32_t partition) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    if (h->stages.empty()) {
        h->stages.resize(h->cfg.n_partitions);
        for (auto &s : h->stages) {
            s.keys.resize(h->cfg.n_keys);
            for (int i = 0; i < h->cfg.n_keys; i++) s.keys[i].type = h->key_widths[i];
            s.payload.resize(h->cfg.n_probe_payload);
            for (int i = 0; i < h->cfg.n_probe_payload; i++) s.payload[i].type = h->probe_payload_types[i];
        }
    }
    return h->stages[partition];
}

void do_probe_push(H *h, int32_t partition, const b200j_batch *b) {
    H::ProbeStage &s = stage_of(h, partition);
    REQUIRE(!h->probe_done[partition], B200J_ERR_STATE, "probe_push after probe_finish");
    check_batch(h, b, h->cfg.n_probe_payload);
    for (int i = 0; i < h->cfg.n_keys; i++)
        REQUIRE(h->key_widths[i] != B200J_T_UTF8, B200J_ERR_UNSUPPORTED, "the coalescing feeder takes fixed-width key columns");
    if (b->n_rows == 0) return;
    REQUIRE(b->n_payload == h->cfg.n_probe_payload, B200J_ERR_INVALID, "probe_push needs every configured probe payload column");
    REQUIRE(s.n + b->n_rows < 0xffff0000LL, B200J_ERR_UNSUPPORTED, "more than 2^32 - 2^16 staged rows: flush first");
    std::vector<void *> temps;
    try {
        // store_append sizes validity bitmaps by build_cap_rows: give it this stage's capacity for the duration of the call
        const int64_t saved = h->build_cap_rows;
        s.cap_rows = std::max<int64_t>(s.cap_rows, std::max<int64_t>(s.n + b->n_rows, (int64_t)1 << 20));
        h->build_cap_rows = s.cap_rows;
        for (int i = 0; i < h->cfg.n_keys; i++) store_append(h, s.keys[i], b->keys[i], s.n, b->n_rows, b->mem, temps);
        for (int i = 0; i < h->cfg.n_probe_payload; i++) store_append(h, s.payload[i], b->payload[i], s.n, b->n_rows, b->mem, temps);
        h->build_cap_rows = saved;
        s.n += b->n_rows;
        CU(cudaStreamSynchronize(h->stream));          // the batch's buffers are borrowed for the call only
        free_temps(h, temps);
    } catch (...) {
        free_temps(h, temps);
        throw;
    }
}

void do_probe_flush(H *h, int32_t partition, uint32_t want, b200j_result *out) {
    H::ProbeStage &s = stage_of(h, partition);
    std::vector<b200j_column> kc(h->cfg.n_keys);
    for (int i = 0; i < h->cfg.n_keys; i++) {
        kc[i].values = s.keys[i].values.p;
        kc[i].offsets = nullptr;
        kc[i].validity = s.keys[i].has_nulls ? (const uint8_t *)s.keys[i].validity.p : nullptr;
        kc[i].bit_offset = 0;
    }
    b200j_batch b{};
    b.struct_size = sizeof(b200j_batch);
    b.mem = B200J_MEM_DEVICE;
    b.n_rows = s.n;
    b.n_keys = h->cfg.n_keys;
    b.n_payload = 0;
    b.keys = kc.data();
    do_probe_batch(h, partition, &b, want, out);
}

void do_gather_staged(H *h, int32_t partition, const b200j_result *pairs, uint32_t want, b200j_gathered *out) {
    H::ProbeStage &s = stage_of(h, partition);
    std::vector<SideCol> sc;
    for (auto *v : {&s.keys, &s.payload})
        for (auto &cs : *v) {
            GatherSrc g{};
            if (s.n > 0) {
                g.values = cs.values.p;
                g.offsets = cs.offsets.p;
                g.offsets64 = 1;
                g.validity = cs.has_nulls ? (const uint8_t *)cs.validity.p : nullptr;
            }
            sc.push_back(SideCol{cs.type, g});
        }
    do_gather_impl(h, pairs, nullptr, want, out, &sc);
    // the stage is consumed: the next push starts a new run of row numbers
    s.n = 0;
    for (auto *v : {&s.keys, &s.payload})
        for (auto &cs : *v) { cs.has_nulls = false; cs.data_bytes = 0; }
}
