// radix_host.inl — host side of the radix-partitioned layout (included inside b200join.cu's anonymous namespace).

void radix_ensure(H *h, int n_bins1, int grid) {
    RadixScratch &rs = h->rscratch;
    const size_t need = (size_t)n_bins1 * (size_t)grid;
    if (rs.hist_cap >= need && rs.bins_cap >= n_bins1) return;
    const size_t cap = std::max(need, rs.hist_cap);
    const int bcap = std::max(n_bins1, rs.bins_cap);
    radix_release(h);
    CU(big_alloc(h, (void **)&rs.cta_hist, cap * 4));
    CU(big_alloc(h, (void **)&rs.bin_tot, (size_t)(bcap + 1) * 4));
    CU(big_alloc(h, (void **)&rs.bin_base, (size_t)(bcap + 1) * 4));
    CU(big_alloc(h, (void **)&rs.off, (size_t)(bcap + 1) * 8));
    CU(cudaHostAlloc((void **)&rs.h_off, (size_t)(bcap + 1) * 8, cudaHostAllocDefault));
    rs.hist_cap = cap;
    rs.bins_cap = bcap;
}

void radix_release(H *h) {
    RadixScratch &rs = h->rscratch;
    if (rs.cta_hist) big_free(h, rs.cta_hist);
    if (rs.bin_tot) big_free(h, rs.bin_tot);
    if (rs.bin_base) big_free(h, rs.bin_base);
    if (rs.off) big_free(h, rs.off);
    if (rs.h_off) cudaFreeHost(rs.h_off);
    rs = RadixScratch{};
}

// Plans the partitioning of `n` keys into n_bins bins (+ the NULL bin): radix mode (n_dest == 0, n_bins = 1 << bits)
// or destination mode (n_bins = n_dest ranks).
PartPlan part_plan(H *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int kw, int64_t n, int bits, int n_dest) {
    REQUIRE(n >= 0 && n < 0xffff0000LL, B200J_ERR_UNSUPPORTED, "more than 2^32 - 2^16 rows in one partitioning pass");
    PartPlan pl;
    pl.kw = kw;
    pl.dest = n_dest > 0;
    PartArgs &a = pl.a;
    a.keys = keys;
    a.validity = validity;
    a.bit_offset = bit_offset;
    a.n = (uint32_t)n;
    a.mis = (uint32_t)(((uintptr_t)keys & 15) / (uintptr_t)kw);
    a.radix_bits = bits;
    a.n_dest = n_dest;
    const bool hot = pl.dest && h->hot_n > 0;          // DEST mode bins: [ranks][hot, kept local][NULL keys]
    a.hot_keys = hot ? (const unsigned long long *)h->hot_table.p : nullptr;
    a.hot_mask = hot ? h->hot_mask : 0u;
    a.n_bins = pl.dest ? (uint32_t)n_dest + (hot ? 1u : 0u) : (1u << bits);
    a.hk = h->hk;
    const int64_t n_tiles = (n + a.mis + SC_TILE - 1) / SC_TILE;
    // DEST mode (multi-GPU exchange): the scatter's stores cross NVLink and its CTAs mostly wait; a smaller grid leaves the
    // SMs' registers and shared memory to the local join running beside it on the other stream (B200J_DEST_CTAS per SM)
    int ctas = SC_CTAS;
    if (pl.dest) {
        const char *e = getenv("B200J_DEST_CTAS");
        ctas = e ? std::max(1, std::min(SC_CTAS, atoi(e))) : SC_CTAS;
    }
    pl.grid = (int)std::max<int64_t>(1, std::min<int64_t>(n_tiles, (int64_t)h->sm_count * ctas));
    a.tiles_per_cta = (uint32_t)((n_tiles + pl.grid - 1) / pl.grid);
    radix_ensure(h, (int)a.n_bins + 1, pl.grid);
    a.cta_hist = h->rscratch.cta_hist;
    a.bin_base = h->rscratch.bin_base;
    return pl;
}

// Histogram + scans. On return (stream synchronised) h->rscratch.h_off[0..n_bins+1] holds the bin offsets
// (h_off[n_bins] = number of non-NULL keys, h_off[n_bins + 1] = n) and, in radix mode, bin_base = those offsets.
void part_hist(H *h, PartPlan &pl, bool sync) {
    RadixScratch &rs = h->rscratch;
    const PartArgs &a = pl.a;
    const int P1 = (int)a.n_bins + 1;
    const size_t hist_smem = (size_t)P1 * 4;
    if (pl.kw == 8) {
        if (pl.dest) part_hist_kernel<8, true><<<pl.grid, RP_THREADS, hist_smem, h->stream>>>(a);
        else part_hist_kernel<8, false><<<pl.grid, RP_THREADS, hist_smem, h->stream>>>(a);
    } else {
        if (pl.dest) part_hist_kernel<4, true><<<pl.grid, RP_THREADS, hist_smem, h->stream>>>(a);
        else part_hist_kernel<4, false><<<pl.grid, RP_THREADS, hist_smem, h->stream>>>(a);
    }
    part_scan_rows_kernel<<<P1, 256, 0, h->stream>>>(rs.cta_hist, (uint32_t)pl.grid, rs.bin_tot);
    part_scan_bins_kernel<<<1, 1024, 0, h->stream>>>(rs.bin_tot, P1, rs.off, pl.dest ? nullptr : rs.bin_base);
    h->st.kernel_launches += 3;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(rs.h_off, rs.off, (size_t)(P1 + 1) * 8, cudaMemcpyDeviceToHost, h->stream));
    if (sync) CU(cudaStreamSynchronize(h->stream));
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute: set for the current device by every b200j_create
// (a process may hold handles on several devices).
void set_kernel_attributes() {
    CU(cudaFuncSetAttribute(part_scatter_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(part_scatter_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(part_scatter_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(part_scatter_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(part_block_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute(part_block_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CU(cudaFuncSetAttribute((part_block_kernel<8, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    if (PB_CLUSTER > 8) {
        CU(cudaFuncSetAttribute((part_block_kernel<8, true>), cudaFuncAttributeNonPortableClusterSizeAllowed, 1));       // clusters of more than 8 CTAs are a non-portable size: opt in
        CU(cudaFuncSetAttribute(part_block_kernel<8>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        CU(cudaFuncSetAttribute(part_block_kernel<4>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    }
}

template <int KW, bool DEST>
void part_scatter_launch(H *h, PartPlan &pl) {
    part_scatter_kernel<KW, DEST><<<pl.grid, SC_THREADS, part_scatter_smem(KW, pl.a.n_bins), h->stream>>>(pl.a);
}

void part_scatter(H *h, PartPlan &pl) {
    if (pl.kw == 8) {
        if (pl.dest) part_scatter_launch<8, true>(h, pl); else part_scatter_launch<8, false>(h, pl);
    } else {
        if (pl.dest) part_scatter_launch<4, true>(h, pl); else part_scatter_launch<4, false>(h, pl);
    }
    h->st.kernel_launches++;
    CU(cudaGetLastError());
}

// Partition `n` keys on the top `bits` hash bits into out_keys / out_rid (both sized for n rows).
// On return the stream is synchronised and h->rscratch.h_off[0..P+1] holds the partition offsets
// (h_off[P] = number of non-NULL keys, h_off[P+1] = n).
void radix_partition(H *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int kw, int64_t n, int bits,
                     void *out_keys, uint32_t *out_rid, double *ms_acc) {
    PhaseTimer tm(h, ms_acc);
    PartPlan pl = part_plan(h, keys, validity, bit_offset, kw, n, bits, 0);
    pl.a.out_keys = out_keys;
    pl.a.out_rid = out_rid;
    pl.a.rid_base = h->probe_rid_base;     // non-zero only while a host batch is probed chunk by chunk
    part_hist(h, pl, false);
    part_scatter(h, pl);
    tm.stop();
    CU(cudaStreamSynchronize(h->stream));
}

int choose_radix_bits(H *h, int64_t n, float lf, int64_t budget) {
    if (h->cfg.radix_bits > 0) return h->cfg.radix_bits;
    // aim at per-partition tables of a quarter of the L2 (half of an explicit l2_budget_bytes): measured on B200, the probe
    // kernel runs at the same speed from 2^7 to 2^10 partitions at 100 M build rows (and 20 % slower at 2^6), while the
    // partition pass gets slower with every bit
    double table_bytes = 16.0 * (double)n / std::max(lf, 0.5f);     // the partition count follows the key count; a lower load factor only spreads a partition
    const double target = std::max<double>(h->cfg.l2_budget_bytes > 0 ? (double)budget / 2.0 : (double)h->l2_bytes / 4.0, 1 << 20);
    int bits = 1;
    while (bits < RP_MAX_BITS && table_bytes / (double)(1 << bits) > target) bits++;
    // the eighth bit is the expensive one: the block-local partition pass takes 7.2 instead of 5.4 ms per 1 B rows and its segments
    // halve to 128 rows. Measured (profiles/r2_o_radix_bits_7_vs_8_crossover.txt): 134 M keys 12.9 (2^7) against 14.6 ms (2^8),
    // 180 M keys 13.7 against 14.9, 250 M keys 15.5 against 15.6 -- 2^7 partitions stay until their tables pass half of the L2
    if (bits == 8 && table_bytes / 128.0 <= 2.0 * target) bits = 7;
    return bits;
}

void radix_build(H *h, const void *keys, const uint8_t *validity, int kw, int64_t n, float lf, int64_t budget) {
    const int bits = choose_radix_bits(h, n, lf, budget);
    const int P = 1 << bits;
    h->radix_bits = bits;
    void *pk = nullptr, *prid = nullptr;
    std::vector<void *> temps;
    try {
        CU(big_alloc(h, &pk, (size_t)n * kw + 16));
        temps.push_back(pk);
        CU(big_alloc(h, &prid, (size_t)n * 4 + 16));
        temps.push_back(prid);
        radix_partition(h, keys, validity, 0, kw, n, bits, pk, (uint32_t *)prid, &h->st.ms_build_partition);
        const unsigned long long *off = h->rscratch.h_off;
        const int64_t n_valid = (int64_t)off[P];

        PhaseTimer tm(h, &h->st.ms_build_insert);
        std::vector<PartEnt> parts(P);
        uint64_t total_buckets = 0;
        for (int p = 0; p < P; p++) {
            const uint64_t cnt = off[p + 1] - off[p];
            const uint64_t nb = table_slots_for((int64_t)cnt, lf) / 2;
            REQUIRE(total_buckets + nb <= 0xffffffffULL, B200J_ERR_UNSUPPORTED, "table too large");
            parts[p].base = (uint32_t)total_buckets;
            parts[p].nb = (uint32_t)nb;
            total_buckets += nb;
        }
        const uint64_t total = total_buckets * 2;
        h->n_slots = total;
        dev_alloc(h, h->slots, (size_t)total * sizeof(Slot));
        const bool fused = h->fused_build && n_valid > 0;
        if (!fused) CU(cudaMemsetAsync(h->slots.p, 0, (size_t)total * sizeof(Slot), h->stream));
        dev_alloc(h, h->parts, (size_t)P * sizeof(PartEnt));
        CU(cudaMemcpyAsync(h->parts.p, parts.data(), (size_t)P * sizeof(PartEnt), cudaMemcpyHostToDevice, h->stream));
        // fused clear + insert (build_insert_fused_kernel): the ordered item sequence C(0), C(1) I(0), ... as prefix sums
        std::vector<uint32_t> meta;
        uint32_t n_items = 0;
        void *d_meta = nullptr;
        if (fused) {
            meta.resize((size_t)(P + 1) + (2 * P + 1) + P + 1, 0u);
            uint32_t *part_off = meta.data(), *seq_off = part_off + P + 1;
            for (int p = 0; p <= P; p++) part_off[p] = (uint32_t)off[p];
            auto clear_items = [&](int p) { return (uint32_t)(((uint64_t)parts[p].nb * 32ull + FB_CLEAR_BYTES - 1) / FB_CLEAR_BYTES); };
            auto insert_items = [&](int p) { return (uint32_t)((off[p + 1] - off[p] + FB_TILE_ROWS - 1) / FB_TILE_ROWS); };
            uint32_t g = 0;
            seq_off[g++] = 0;
            n_items = clear_items(0);
            for (int p = 0; p + 1 < P; p++) {
                seq_off[g++] = n_items; n_items += clear_items(p + 1);
                seq_off[g++] = n_items; n_items += insert_items(p);
            }
            seq_off[g++] = n_items; n_items += insert_items(P - 1);
            seq_off[g] = n_items;                         // g == 2P
            CU(big_alloc(h, &d_meta, meta.size() * 4 + 16));
            temps.push_back(d_meta);
            CU(cudaMemcpyAsync(d_meta, meta.data(), meta.size() * 4, cudaMemcpyHostToDevice, h->stream));
        }
        CU(cudaStreamSynchronize(h->stream));      // parts / meta are stack vectors

        BuildArgs a{};
        a.hk = h->hk;
        a.ctr = h->d_ctr;
        a.keys = pk;
        a.validity = nullptr;
        a.rowid = (const uint32_t *)prid;
        a.n = n_valid;
        a.mis = 0;
        a.tv = h->view();
        if (n_valid > 0 && total < 0xffffffffULL) {          // slot indices fit 32 bits: remember them for the duplicate-run pass
            void *so = nullptr;
            CU(big_alloc(h, &so, (size_t)n_valid * 4 + 16));
            temps.push_back(so);
            a.slot_of = (uint32_t *)so;
            a.dup_bit = (!fused && total < 0x7fffffffULL) ? 1u : 0u;      // build_insert_kernel marks the rows csr_fill has to place
        }
        if (fused) {
            FusedBuildArgs f{};
            f.b = a;
            f.part_off = (const uint32_t *)d_meta;
            f.seq_off = f.part_off + P + 1;
            f.n_parts = (uint32_t)P;
            f.n_items = n_items;
            f.done = (unsigned int *)d_meta + (P + 1) + (2 * P + 1);
            f.error = f.done + P;
            const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)n_items, (int64_t)h->sm_count * h->fused_ctas_per_sm));
            launch_by_kw(kw, [&](auto KW) { build_insert_fused_kernel<decltype(KW)::value><<<grid, TILE_THREADS, 0, h->stream>>>(f); });
            h->st.kernel_launches++;
            CU(cudaGetLastError());
            uint32_t err = 0;
            CU(cudaMemcpyAsync(&err, f.error, 4, cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
            REQUIRE(err == 0, B200J_ERR_CUDA, "fused build: a clear item did not finish in time (watchdog)");
        } else if (n_valid > 0) {
            const int grid = grid_for(h, n_valid, TILE_ROWS, 8);
            launch_by_kw(kw, [&](auto KW) { build_insert_kernel<decltype(KW)::value, true><<<grid, TILE_THREADS, 0, h->stream>>>(a); });
            h->st.kernel_launches++;
        }
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        h->h_ctr->null_keys = (unsigned long long)(n - n_valid);
        if (h->h_ctr->dup_rows > 0) {
            const uint64_t need = h->h_ctr->dup_rows + 2 * h->h_ctr->multi_slots + 64;
            REQUIRE(need < 0xffffffffULL, B200J_ERR_UNSUPPORTED, "duplicate-run index exceeds 32 bits");
            dev_alloc(h, h->rowids, (size_t)need * 4);
            a.tv = h->view();
            csr_scan_kernel<<<grid_for(h, (int64_t)total + 1, 256, 8), 256, 0, h->stream>>>(a.tv, total, h->d_ctr);
            const int grid = grid_for(h, n_valid, TILE_ROWS, 8);
            launch_by_kw(kw, [&](auto KW) { csr_fill_kernel<decltype(KW)::value, true><<<grid, TILE_THREADS, 0, h->stream>>>(a); });
            h->st.kernel_launches += 2;
            CU(cudaGetLastError());
        }
        tm.stop();
        free_temps(h, temps);
    } catch (...) {
        free_temps(h, temps);
        throw;
    }
}

// ---- multi-GPU exchange: count, then scatter into per-destination buffers (local or peer memory) ----------------
void do_shuffle_count(H *h, const b200j_batch *b, int32_t n_dest, int64_t *counts) {
    check_batch(h, b, 0);
    REQUIRE(h->cfg.n_keys == 1, B200J_ERR_UNSUPPORTED, "the exchange takes a single 4/8-byte key column");
    REQUIRE(n_dest >= 1 && n_dest <= B200J_MAX_DEST, B200J_ERR_INVALID, "n_dest must be in 1..16");
    REQUIRE(counts != nullptr, B200J_ERR_INVALID, "counts is NULL");
    REQUIRE(b->mem == B200J_MEM_DEVICE || b->n_rows == 0, B200J_ERR_UNSUPPORTED, "the exchange takes device-resident batches");
    REQUIRE(h->cfg.n_keys == 1 && h->key_widths[0] != B200J_T_UTF8, B200J_ERR_UNSUPPORTED, "the exchange partitions one fixed-width key column");
    h->shuffle_armed = false;
    const bool hot = h->hot_n > 0;
    for (int d = 0; d <= n_dest + (hot ? 1 : 0); d++) counts[d] = 0;
    if (b->n_rows == 0) {
        h->shuffle_n = 0;
        h->shuffle_armed = true;
        return;
    }
    const b200j_column &kc = b->keys[0];
    h->shuffle_plan = part_plan(h, kc.values, kc.validity, kc.bit_offset, h->key_widths[0], b->n_rows, 0, n_dest);
    PhaseTimer tm(h, &h->st.ms_exchange_partition);
    part_hist(h, h->shuffle_plan, true);
    tm.stop();
    const unsigned long long *off = h->rscratch.h_off;
    for (int d = 0; d < n_dest; d++) counts[d] = (int64_t)(off[d + 1] - off[d]);
    if (hot) {      // internal bin order [ranks][hot][NULL] -> API order [ranks][NULL][hot]
        counts[n_dest + 1] = (int64_t)(off[n_dest + 1] - off[n_dest]);
        counts[n_dest] = (int64_t)(off[n_dest + 2] - off[n_dest + 1]);
    } else {
        counts[n_dest] = (int64_t)(off[n_dest + 1] - off[n_dest]);
    }
    h->shuffle_n = b->n_rows;
    h->shuffle_armed = true;
}

void do_shuffle_scatter(H *h, int32_t n_dest, void *const *dest_keys, uint32_t *const *dest_rowid, const int64_t *dest_offset, uint32_t rowid_base) {
    REQUIRE(h->shuffle_armed, B200J_ERR_STATE, "b200j_shuffle_scatter without a preceding b200j_shuffle_count");
    h->shuffle_armed = false;
    if (h->shuffle_n == 0) return;
    PartPlan &pl = h->shuffle_plan;
    REQUIRE(n_dest == pl.a.n_dest, B200J_ERR_INVALID, "n_dest differs from the b200j_shuffle_count call");
    REQUIRE(dest_keys && dest_rowid && dest_offset, B200J_ERR_INVALID, "destination arrays missing");
    uint32_t base[B200J_MAX_DEST + 2];
    const unsigned long long *off = h->rscratch.h_off;
    const bool hot = pl.a.hot_keys != nullptr;
    REQUIRE(hot == (h->hot_n > 0), B200J_ERR_STATE, "the hot-key set changed between b200j_shuffle_count and b200j_shuffle_scatter");
    const int n_api = n_dest + 1 + (hot ? 1 : 0);
    for (int d = 0; d < n_api; d++) {
        // API order [ranks][NULL][hot] -> internal bin order [ranks][hot][NULL]
        const int bin = d < n_dest ? d : (hot ? (d == n_dest ? n_dest + 1 : n_dest) : n_dest);
        const bool used = off[bin + 1] > off[bin];
        REQUIRE(!used || (dest_keys[d] && dest_rowid[d]), B200J_ERR_INVALID, "destination buffer missing for a non-empty destination");
        REQUIRE(dest_offset[d] >= 0 && dest_offset[d] < 0xffff0000LL, B200J_ERR_INVALID, "destination offset out of range");
        base[bin] = (uint32_t)dest_offset[d];
        pl.a.dest_keys[bin] = dest_keys[d];
        pl.a.dest_rid[bin] = dest_rowid[d];
    }
    pl.a.rid_base = rowid_base;
    PhaseTimer tm(h, &h->st.ms_exchange_partition);
    CU(cudaMemcpyAsync(h->rscratch.bin_base, base, (size_t)n_api * 4, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));      // `base` is a stack array
    part_scatter(h, pl);
    tm.stop();
}

// Single-pass local scatter by destination (b200j_shuffle_local_scatter): regions of `region_rows` rows per bin, reserved tile
// by tile with atomics; counts come back in the API order [ranks][NULL][hot].
void do_shuffle_local_scatter(H *h, const b200j_batch *b, int32_t n_dest, void *send_keys, uint32_t *send_rowid, int64_t region_rows,
                              uint32_t rowid_base, int32_t key_out_width, int64_t *counts) {
    check_batch(h, b, 0);
    REQUIRE(h->cfg.n_keys == 1 && h->key_widths[0] != B200J_T_UTF8, B200J_ERR_UNSUPPORTED, "the exchange partitions one fixed-width key column");
    REQUIRE(n_dest >= 1 && n_dest <= B200J_MAX_DEST, B200J_ERR_INVALID, "n_dest must be in 1..16");
    REQUIRE(counts != nullptr, B200J_ERR_INVALID, "counts is NULL");
    REQUIRE(b->mem == B200J_MEM_DEVICE || b->n_rows == 0, B200J_ERR_UNSUPPORTED, "the exchange takes device-resident batches");
    REQUIRE(region_rows > 0 && region_rows < 0xffff0000LL, B200J_ERR_INVALID, "region_rows out of range");
    REQUIRE(!h->shuffle_armed, B200J_ERR_STATE, "b200j_shuffle_local_scatter between b200j_shuffle_count and b200j_shuffle_scatter");
    const bool hot = h->hot_n > 0;
    const int n_api = n_dest + 1 + (hot ? 1 : 0);
    for (int d = 0; d < n_api; d++) counts[d] = 0;
    if (b->n_rows == 0) return;
    REQUIRE(send_keys && send_rowid, B200J_ERR_INVALID, "send buffers missing");
    const int kw = h->key_widths[0];
    const int okw = key_out_width > 0 ? key_out_width : kw;
    REQUIRE(okw == kw || (okw == 4 && kw == 8), B200J_ERR_INVALID, "key_out_width must be the key width, or 4 for 8-byte keys");
    const b200j_column &kc = b->keys[0];
    PartPlan pl = part_plan(h, kc.values, kc.validity, kc.bit_offset, kw, b->n_rows, 0, n_dest);
    if (!h->d_reserve) CU(cudaMalloc((void **)&h->d_reserve, (B200J_MAX_DEST + 4) * 4));
    CU(cudaMemsetAsync(h->d_reserve, 0, (B200J_MAX_DEST + 4) * 4, h->stream));
    pl.a.reserve = h->d_reserve;
    pl.a.reserve_cap = (uint32_t)region_rows;
    pl.a.reserve_ovf = h->d_reserve + B200J_MAX_DEST + 3;
    pl.a.narrow_out = okw != kw ? 1u : 0u;
    pl.a.narrow_ovf = h->d_reserve + B200J_MAX_DEST + 2;
    for (int d = 0; d < n_api; d++) {
        // API order [ranks][NULL][hot] -> internal bin order [ranks][hot][NULL]
        const int bin = d < n_dest ? d : (hot ? (d == n_dest ? n_dest + 1 : n_dest) : n_dest);
        pl.a.dest_keys[bin] = (char *)send_keys + (size_t)d * (size_t)region_rows * okw;
        pl.a.dest_rid[bin] = send_rowid + (size_t)d * (size_t)region_rows;
    }
    pl.a.rid_base = rowid_base;
    {
        PhaseTimer tm(h, &h->st.ms_exchange_partition);
        part_scatter(h, pl);
        tm.stop();
    }
    uint32_t got[B200J_MAX_DEST + 4];
    CU(cudaMemcpyAsync(got, h->d_reserve, sizeof(got), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    REQUIRE(got[B200J_MAX_DEST + 3] == 0, B200J_ERR_INVALID, "region too small: a destination received more rows than region_rows");
    REQUIRE(okw == kw || got[B200J_MAX_DEST + 2] == 0, B200J_ERR_INVALID, "keys do not fit 32 bits: repeat with key_out_width = 8");
    for (int d = 0; d < n_api; d++) {
        const int bin = d < n_dest ? d : (hot ? (d == n_dest ? n_dest + 1 : n_dest) : n_dest);
        counts[d] = (int64_t)got[bin];
    }
}

// Heavy-hitter key set of the exchange: rows with these keys stay on the rank that holds them (see part_is_hot).
void do_shuffle_set_hot(H *h, const uint64_t *keys, int32_t n) {
    REQUIRE(n >= 0 && n <= 65536, B200J_ERR_INVALID, "hot key count must be in 0..65536");
    REQUIRE(n == 0 || keys != nullptr, B200J_ERR_INVALID, "hot keys missing");
    REQUIRE(!h->shuffle_armed, B200J_ERR_STATE, "b200j_shuffle_set_hot between b200j_shuffle_count and b200j_shuffle_scatter");
    h->hot_n = 0;
    if (n == 0) return;
    uint32_t size = 64;
    while (size < (uint32_t)n * 4u) size <<= 1;
    std::vector<unsigned long long> tab(size, 0ULL);
    int kept = 0;
    for (int i = 0; i < n; i++) {
        const unsigned long long k = keys[i];
        if (k == 0ULL) continue;                       // the key value 0 marks an empty slot: it is never routed as hot
        uint32_t j = (uint32_t)((k * 0x9E3779B97F4A7C15ULL) >> 32) & (size - 1);
        while (tab[j] != 0ULL && tab[j] != k) j = (j + 1) & (size - 1);
        if (tab[j] == 0ULL) { tab[j] = k; kept++; }
    }
    if (kept == 0) return;
    dev_alloc(h, h->hot_table, (size_t)size * 8);
    CU(cudaMemcpyAsync(h->hot_table.p, tab.data(), (size_t)size * 8, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));              // `tab` is a stack-owned vector
    h->hot_mask = size - 1;
    h->hot_n = kept;
}


// ---- block-local partition + segment-walking dense probe (radix_block.cuh, probe_dense_seg_kernel) -----------------------
template <int KW, bool NARROW = false>
void part_block_launch(H *h, const BlockPartArgs &a, uint32_t n_bins) {
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(PB_THREADS);
    cfg.dynamicSmemBytes = part_block_smem(KW, n_bins);
    cfg.stream = h->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PB_CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // clusters that can be resident at once (2 CTAs per SM, a cluster inside one GPC): a persistent grid of that many
    int &cached = h->pb_clusters[KW == 8 ? 1 : 0][a.radix_bits];
    if (cached == 0) {
        cfg.gridDim = dim3((unsigned)(h->sm_count * PB_CTAS / PB_CLUSTER * PB_CLUSTER));
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, part_block_kernel<KW, NARROW>, &cfg) != cudaSuccess || n <= 0) {
            cudaGetLastError();
            n = std::max(1, h->sm_count * PB_CTAS / PB_CLUSTER);
        }
        cached = n;
    }
    const int n_clusters = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)a.n_blocks, (int64_t)cached));
    cfg.gridDim = dim3((unsigned)(n_clusters * PB_CLUSTER));
    CU(cudaLaunchKernelEx(&cfg, part_block_kernel<KW, NARROW>, a));
    h->st.kernel_launches++;
}

// One piece of a probe batch: `n` device-resident keys (possibly in a peer GPU's memory mapped with b200j_ipc_open: the
// partition pass then PULLS them over NVLink while it partitions -- the multi-GPU exchange fused into the local join).
struct ProbePiece {
    const void    *keys;
    const uint8_t *validity;
    int64_t        bit_offset;
    int64_t        n;
};

// Dense probe of device-resident keys against unique build keys behind the radix layout; the batch is the concatenation
// of `pieces` (probe indices count across the pieces). On return ro owns the result arrays; returns the number of result
// rows. PM_OUTER keeps the rows without a match (NULL build index), PM_PAIRS pays one compaction pass when there were any
// (NULL-key rows count as rows without a match).
int64_t probe_dense_blocked_pieces(H *h, const ProbePiece *pieces, int n_pieces, const ProbePlan &pl, ResultOwner *ro, int64_t *n_build_null,
                                   int piece_kw) {
    const int kw = piece_kw > 0 ? piece_kw : h->table_kw;      // key width of the pieces: the table's, or 4 = zero-extended values of 8-byte keys
    const uint32_t P = 1u << h->radix_bits;
    int64_t n = 0;
    uint32_t n_blocks = 0;
    std::vector<uint32_t> piece_blocks(n_pieces), piece_mis(n_pieces);
    for (int i = 0; i < n_pieces; i++) {
        piece_mis[i] = (uint32_t)(((uintptr_t)pieces[i].keys & 15) / (uintptr_t)kw);
        const uint32_t n_tiles = (uint32_t)(((uint64_t)pieces[i].n + piece_mis[i] + PB_TILE - 1) / PB_TILE);
        piece_blocks[i] = pieces[i].n > 0 ? (n_tiles + PB_CLUSTER - 1) / PB_CLUSTER : 0u;
        n_blocks += piece_blocks[i];
        n += pieces[i].n;
    }
    REQUIRE(n < 0xffff0000LL, B200J_ERR_UNSUPPORTED, "probe batch of more than 2^32 - 2^16 rows (probe indices are UInt32)");
    void *tk = nullptr, *tr = nullptr, *seg = nullptr, *ob = nullptr, *ob2 = nullptr, *op2 = nullptr, *oslot = nullptr;
    // B200J_OUT_GATHER_HINT: the probe also writes the table slot of every row (slot indices must fit 32 bits) -- when there is a
    // fixed-width build payload column for the gather to read through it; a join without one keeps only the key stream
    bool fixed_payload = false;
    for (int i = 0; i < h->cfg.n_build_payload; i++) fixed_payload |= h->bpayload[i].type != B200J_T_UTF8;
    // ... and when the probe side is big enough to pay for the slot-ordered copies (a pass over the whole table, ~3 ms per 100 M
    // build rows and column, against ~12 ms saved per 1 B gathered rows and column): at least two probe rows per build row so far
    // ... and not when the last batch of an Inner / Left join had probe rows without a match: such a result is compacted and
    // leaves the stream's order, so nothing of this could be used
    const bool want_slot = ro->want_aux && ro->want_seg && fixed_payload && h->n_slots + 2 < 0xffffffffULL &&
                           h->probed_since_build + n >= 2 * h->n_build && !(pl.mode == PM_PAIRS && h->miss_frac > 0.f);
    h->probed_since_build += n;
    Trace tr_;
    try {
        // 8-byte keys travel through the partitioned stream as 32-bit values while they fit (see part_block_kernel NARROW)
        bool narrow = kw == 8 && h->narrow_ok;
        uint32_t *d_ovf = (uint32_t *)(h->d_part_rows + (1 << PB_MAX_BITS) + 1);
    again:
        const int skw = narrow ? 4 : kw;                 // key width in the partitioned stream
        if (tk) { big_free(h, tk); tk = nullptr; }
        CU(big_alloc(h, &tk, (size_t)n * skw + 16));
        if (!tr) CU(big_alloc(h, &tr, (size_t)n * 4 + 16));
        if (!seg) CU(big_alloc(h, &seg, (size_t)n_blocks * (P + 3) * 4 + 16));       // [n_blocks][P + 2] segment bounds, then [n_blocks] block bases
        if (!ob) CU(big_alloc(h, &ob, (size_t)n * 8 + 16));
        if (want_slot && !oslot) CU(big_alloc(h, &oslot, (size_t)n * 4 + 16));
        uint32_t *block_base = (uint32_t *)seg + (size_t)n_blocks * (P + 2);
        CU(cudaMemsetAsync(h->d_part_rows, 0, ((size_t)(1 << PB_MAX_BITS) + 2) * 8, h->stream));
        {
            PhaseTimer tm(h, &h->st.ms_probe_partition);
            uint32_t blk0 = 0;
            int64_t row0 = 0;
            for (int i = 0; i < n_pieces; i++) {
                if (pieces[i].n == 0) continue;
                BlockPartArgs ba{};
                ba.keys = pieces[i].keys;
                ba.validity = pieces[i].validity;
                ba.bit_offset = pieces[i].bit_offset;
                ba.n = (uint32_t)pieces[i].n;
                ba.mis = piece_mis[i];
                ba.radix_bits = h->radix_bits;
                ba.n_blocks = piece_blocks[i];
                ba.hk = h->hk;
                ba.out_keys = tk;
                ba.out_rid = (uint32_t *)tr;
                ba.seg = (uint32_t *)seg + (size_t)blk0 * (P + 2);
                ba.block_base = block_base + blk0;
                ba.part_rows = h->d_part_rows;
                ba.rid_base = h->probe_rid_base;
                ba.out_base = (uint32_t)row0;
                ba.narrow_ovf = d_ovf;
                if (narrow) part_block_launch<8, true>(h, ba, P);
                else if (kw == 8) part_block_launch<8>(h, ba, P);
                else part_block_launch<4>(h, ba, P);
                CU(cudaGetLastError());
                blk0 += piece_blocks[i];
                row0 += pieces[i].n;
            }
            tm.stop();
        }
        CU(cudaMemcpyAsync(h->h_part_rows, h->d_part_rows, ((size_t)(1 << PB_MAX_BITS) + 2) * 8, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        tr_.mark("block partition");
        if (narrow && (uint32_t)h->h_part_rows[(1 << PB_MAX_BITS) + 1] != 0u) {
            // a key outside [0, 2^32): once more with 8-byte keys, and no further attempts for this join
            narrow = false;
            h->narrow_ok = false;
            goto again;
        }
        const int64_t n_null = (int64_t)h->h_part_rows[P];
        int64_t biggest = 0;
        for (uint32_t p = 0; p < P; p++) biggest = std::max<int64_t>(biggest, (int64_t)h->h_part_rows[p]);
        const bool l1 = h->skew_l1 && P >= 8 && biggest * (int64_t)P > 3 * (n - n_null);

        ProbeArgs a{};
        a.keys = tk;
        a.n = n;
        a.tv = h->view();
        a.hk = h->hk;
        a.out_build = (uint64_t *)ob;
        a.out_probe = (uint32_t *)tr;
        a.out_cap = (uint64_t)n;
        a.visited = (uint32_t *)h->visited.p;
        a.ctr = h->d_ctr;
        SegArgs sg{};
        sg.seg = (const uint32_t *)seg;
        sg.block_base = block_base;
        sg.n_blocks = n_blocks;
        sg.n_units = n_blocks * (n_null > 0 ? P + 1 : P);
        sg.out_slot = (uint32_t *)oslot;
        sg.zero_slot_index = (uint32_t)h->n_slots;
        {
            PhaseTimer tm(h, &h->st.ms_probe);
            const int64_t n_claims = ((int64_t)sg.n_units + SEG_UNITS - 1) / SEG_UNITS;
            const int grid = grid_for(h, n_claims, PD_WARPS, h->seg_ctas_per_sm);
            auto launch = [&](auto KW, auto MARK) {
                constexpr int kw_ = decltype(KW)::value;
                constexpr bool mk_ = decltype(MARK)::value;
                if (want_slot) {
                    if (l1) probe_dense_seg_kernel<kw_, mk_, true, true><<<grid, PD_THREADS, 0, h->stream>>>(a, sg);
                    else probe_dense_seg_kernel<kw_, mk_, false, true><<<grid, PD_THREADS, 0, h->stream>>>(a, sg);
                } else if (l1) probe_dense_seg_kernel<kw_, mk_, true><<<grid, PD_THREADS, 0, h->stream>>>(a, sg);
                else probe_dense_seg_kernel<kw_, mk_, false><<<grid, PD_THREADS, 0, h->stream>>>(a, sg);
            };
            if (skw == 8) {
                if (pl.mark) launch(std::integral_constant<int, 8>{}, std::true_type{});
                else launch(std::integral_constant<int, 8>{}, std::false_type{});
            } else {
                if (pl.mark) launch(std::integral_constant<int, 4>{}, std::true_type{});
                else launch(std::integral_constant<int, 4>{}, std::false_type{});
            }
            h->st.kernel_launches++;
            h->st.probe_launches++;
            h->st.dense_probe_launches++;
            h->st.seg_probe_launches++;
            if (narrow) h->st.narrow_probe_launches++;
            CU(cudaGetLastError());
            tm.stop();
        }
        CU(cudaMemcpyAsync(h->h_ctr, h->d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        tr_.mark("segment probe + counters");
        const int64_t misses = (int64_t)h->h_ctr->build_null_out;       // rows without a match, NULL-key rows included
        h->miss_frac = n > n_null ? (float)((double)(misses - n_null) / (double)(n - n_null)) : 0.f;
        if (pl.mode == PM_OUTER || misses == 0) {
            *n_build_null = misses;
            ro->build_idx = ob;
            ro->probe_idx = tr;
            if (ro->want_seg && h->gather_seg) {
                // the result keeps the stream's order: its segment table goes with it (b200j_gather walks the build side by partition)
                ro->seg = seg;
                ro->seg_blocks = n_blocks;
                ro->seg_P = P;
                ro->seg_n = n;
                ro->aux_misses = misses;
                h->seg_owners.insert(ro);
                if (ro->want_aux) {
                    // ... and with B200J_OUT_GATHER_HINT the partitioned keys (= the probe key of every result row; not when the
                    // batch has NULL keys, whose rows carry no key) and the slot of every pair
                    if (n_null == 0) { ro->aux_keys = tk; ro->aux_kw = skw; tk = nullptr; }
                    ro->aux_slot = oslot;
                    oslot = nullptr;
                }
            } else {
                big_free(h, seg);
            }
            seg = nullptr;
            if (tk) big_free(h, tk);
            if (oslot) big_free(h, oslot);
            tk = oslot = nullptr;
            return n;
        }
        big_free(h, tk); tk = nullptr;
        big_free(h, seg); seg = nullptr;
        if (oslot) { big_free(h, oslot); oslot = nullptr; }
        const int64_t keep = n - misses;
        CU(big_alloc(h, &ob2, (size_t)std::max<int64_t>(keep, 1) * 8));
        CU(big_alloc(h, &op2, (size_t)std::max<int64_t>(keep, 1) * 4));
        CU(cudaMemsetAsync(h->d_ctr, 0, COUNTERS_LAUNCH_BYTES, h->stream));
        {
            PhaseTimer tm(h, &h->st.ms_probe);
            compact_pairs_kernel<<<grid_for(h, n, CP_TILE, 8), CP_THREADS, 0, h->stream>>>((const uint64_t *)ob, (const uint32_t *)tr, (uint64_t)n,
                                                                                           (uint64_t *)ob2, (uint32_t *)op2, &h->d_ctr->out_count);
            h->st.kernel_launches++;
            h->st.compact_launches++;
            CU(cudaGetLastError());
            tm.stop();
        }
        CU(cudaStreamSynchronize(h->stream));
        big_free(h, ob);
        big_free(h, tr);
        *n_build_null = 0;
        ro->build_idx = ob2;
        ro->probe_idx = op2;
        return keep;
    } catch (...) {
        for (void *p : {tk, tr, seg, ob, ob2, op2, oslot})
            if (p) big_free(h, p);
        throw;
    }
}

int64_t probe_dense_blocked(H *h, const void *d_keys, const uint8_t *d_validity, int64_t bit_offset, int64_t n, const ProbePlan &pl,
                            ResultOwner *ro, int64_t *n_build_null) {
    const ProbePiece one{d_keys, d_validity, bit_offset, n};
    return probe_dense_blocked_pieces(h, &one, 1, pl, ro, n_build_null);
}

// b200j_probe_pieces: see include/b200join.h. The dense path behind the radix layout takes the pieces as they are; every
// other path gets them concatenated into one device buffer first.
void do_probe_pieces(H *h, int32_t partition, int32_t n_pieces, const void *const *piece_keys, const int64_t *piece_rows, int32_t piece_key_width,
                     uint32_t want, b200j_result *out) {
    REQUIRE(partition >= 0 && partition < h->cfg.n_partitions, B200J_ERR_INVALID, "partition out of range");
    REQUIRE(h->built, B200J_ERR_INVALID, "probe before every partition finished the build side");
    REQUIRE(!h->probe_done[partition], B200J_ERR_STATE, "probe_pieces after probe_finish");
    REQUIRE(!h->shuffle_armed, B200J_ERR_STATE, "probe_pieces between b200j_shuffle_count and b200j_shuffle_scatter");
    REQUIRE(out != nullptr && out->struct_size >= RESULT_V1_SIZE, B200J_ERR_INVALID, "result struct missing or too small");
    REQUIRE(!h->composite, B200J_ERR_UNSUPPORTED, "probe_pieces takes a single fixed-width key column");
    REQUIRE(n_pieces >= 0 && n_pieces <= 64, B200J_ERR_INVALID, "n_pieces must be in 0..64");
    REQUIRE(n_pieces == 0 || (piece_keys && piece_rows), B200J_ERR_INVALID, "piece arrays missing");
    const ProbePlan pl = plan_for(h->cfg.join_type, (want & B200J_OUT_MATCHES_ONLY) != 0);
    const bool count_only = (want & B200J_OUT_COUNT_ONLY) != 0;
    const int kw = h->key_widths[0];
    const int pkw = piece_key_width > 0 ? piece_key_width : kw;
    REQUIRE(pkw == kw || (pkw == 4 && kw == 8), B200J_ERR_INVALID, "piece_key_width must be the key width, or 4 for 8-byte keys");
    std::vector<ProbePiece> pieces;
    int64_t n = 0;
    for (int i = 0; i < n_pieces; i++) {
        REQUIRE(piece_rows[i] >= 0, B200J_ERR_INVALID, "negative piece length");
        REQUIRE(piece_rows[i] == 0 || piece_keys[i] != nullptr, B200J_ERR_INVALID, "piece without keys");
        pieces.push_back(ProbePiece{piece_keys[i], nullptr, 0, piece_rows[i]});
        n += piece_rows[i];
    }
    REQUIRE(n < 0xffff0000LL, B200J_ERR_UNSUPPORTED, "probe batch of more than 2^32 - 2^16 rows (probe indices are UInt32)");
    ResultOwner *ro = new ResultOwner();
    ro->want_seg = true;
    ro->want_aux = ((want & B200J_OUT_GATHER_HINT) != 0 || (h->auto_hint && h->gather_seen)) && !(want & B200J_OUT_HOST) && !count_only && h->gather_seg;
    void *cat = nullptr;
    try {
        int64_t total = 0, n_bnull = 0;
        const bool blocked = h->layout == B200J_LAYOUT_RADIX && !count_only && h->dense_ok && h->block_part_ok && h->radix_bits <= PB_MAX_BITS &&
                             h->st.duplicate_rows == 0 && (pl.mode == PM_PAIRS || pl.mode == PM_OUTER);
        if (n == 0) {
            total = 0;
        } else if (blocked) {
            total = probe_dense_blocked_pieces(h, pieces.data(), (int)pieces.size(), pl, ro, &n_bnull, pkw);
        } else {
            CU(big_alloc(h, &cat, (size_t)n * kw + 16));
            int64_t at = 0;
            for (const ProbePiece &pc : pieces) {
                if (pc.n && pkw == kw) CU(cudaMemcpyAsync((char *)cat + (size_t)at * kw, pc.keys, (size_t)pc.n * kw, cudaMemcpyDefault, h->stream));
                else if (pc.n) {
                    widen_u32_u64_kernel<<<grid_for(h, pc.n, 256, 8), 256, 0, h->stream>>>((const uint32_t *)pc.keys, (unsigned long long *)cat + at, pc.n);
                    h->st.kernel_launches++;
                    CU(cudaGetLastError());
                }
                at += pc.n;
            }
            total = probe_device(h, cat, nullptr, 0, n, pl, count_only, ro, &n_bnull);
        }
        if (count_only) {
            if (ro->build_idx) big_free(h, ro->build_idx);
            if (ro->probe_idx) big_free(h, ro->probe_idx);
            ro->build_idx = ro->probe_idx = nullptr;
        } else {
            if (want & B200J_OUT_BUILD_U32) narrow_result(h, ro, total);
            if ((want & B200J_OUT_HOST) && ro->mem != B200J_MEM_HOST) move_result_to_host(h, ro, total);
            else if (want & B200J_OUT_BUILD_U32) CU(cudaStreamSynchronize(h->stream));
        }
        if (cat) big_free(h, cat);
        cat = nullptr;
        fill_result(out, ro, total, ro->mem);
        out->n_build_null = n_bnull;
        out->n_probe_null = 0;
        h->st.probe_rows += n;
        h->st.result_rows += total;
    } catch (...) {
        if (cat) big_free(h, cat);
        release_owner(h, ro);
        throw;
    }
}
