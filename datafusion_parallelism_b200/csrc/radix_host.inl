// radix_host.inl — host side of the radix-partitioned layout (included inside b200join.cu's anonymous namespace).

void radix_ensure(H *h, int bits) {
    RadixScratch &rs = h->rscratch;
    if (rs.alloc_bits >= bits) return;
    if (rs.hist) cudaFreeAsync(rs.hist, h->stream);
    if (rs.off) cudaFreeAsync(rs.off, h->stream);
    if (rs.cursor) cudaFreeAsync(rs.cursor, h->stream);
    if (rs.h_off) cudaFreeHost(rs.h_off);
    rs = RadixScratch{};
    const size_t bytes = (((size_t)1 << bits) + 2) * 8;
    CU(cudaMallocAsync((void **)&rs.hist, bytes, h->stream));
    CU(cudaMallocAsync((void **)&rs.off, bytes, h->stream));
    CU(cudaMallocAsync((void **)&rs.cursor, bytes, h->stream));
    CU(cudaHostAlloc((void **)&rs.h_off, bytes, cudaHostAllocDefault));
    rs.alloc_bits = bits;
}

void radix_release(H *h) {
    RadixScratch &rs = h->rscratch;
    if (rs.hist) cudaFreeAsync(rs.hist, h->stream);
    if (rs.off) cudaFreeAsync(rs.off, h->stream);
    if (rs.cursor) cudaFreeAsync(rs.cursor, h->stream);
    if (rs.h_off) cudaFreeHost(rs.h_off);
    rs = RadixScratch{};
}

// Partition `n` keys on the top `bits` hash bits into out_keys / out_rid (both sized for n rows).
// On return the stream is synchronised and h->rscratch.h_off[0..P+1] holds the partition offsets
// (h_off[P] = number of non-NULL keys, h_off[P+1] = n).
void radix_partition(H *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int kw, int64_t n, int bits,
                     void *out_keys, uint32_t *out_rid, double *ms_acc) {
    radix_ensure(h, bits);
    RadixScratch &rs = h->rscratch;
    const int P1 = (1 << bits) + 1;
    PhaseTimer tm(h, ms_acc);
    CU(cudaMemsetAsync(rs.hist, 0, (size_t)(P1 + 1) * 8, h->stream));
    PartArgs a{};
    a.keys = keys;
    a.validity = validity;
    a.bit_offset = bit_offset;
    a.n = n;
    a.mis = (int)(((uintptr_t)keys & 15) / (uintptr_t)kw);
    a.radix_bits = bits;
    a.hk = h->hk;
    a.hist = rs.hist;
    a.cursor = rs.cursor;
    a.out_keys = out_keys;
    a.out_rid = out_rid;
    const int grid = grid_for(h, n + a.mis, RP_TILE, 2);
    const size_t hist_smem = (size_t)P1 * 4;
    const size_t sc_smem = radix_scatter_smem(kw, bits);
    if (kw == 8) {
        static bool attr8 = false;
        if (!attr8) { CU(cudaFuncSetAttribute(radix_scatter_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); attr8 = true; }
        radix_hist_kernel<8><<<grid, RP_THREADS, hist_smem, h->stream>>>(a);
        radix_scan_kernel<<<1, 1024, 0, h->stream>>>(rs.hist, P1, rs.off, rs.cursor);
        radix_scatter_kernel<8><<<grid, RP_THREADS, sc_smem, h->stream>>>(a);
    } else {
        static bool attr4 = false;
        if (!attr4) { CU(cudaFuncSetAttribute(radix_scatter_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); attr4 = true; }
        radix_hist_kernel<4><<<grid, RP_THREADS, hist_smem, h->stream>>>(a);
        radix_scan_kernel<<<1, 1024, 0, h->stream>>>(rs.hist, P1, rs.off, rs.cursor);
        radix_scatter_kernel<4><<<grid, RP_THREADS, sc_smem, h->stream>>>(a);
    }
    h->st.kernel_launches += 3;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(rs.h_off, rs.off, (size_t)(P1 + 1) * 8, cudaMemcpyDeviceToHost, h->stream));
    tm.stop();
    CU(cudaStreamSynchronize(h->stream));
}

int choose_radix_bits(H *h, int64_t n, float lf, int64_t budget) {
    if (h->cfg.radix_bits > 0) return h->cfg.radix_bits;
    // aim at per-partition tables of a quarter of the L2 budget: two or three of them are live at any time
    double table_bytes = 16.0 * (double)n / lf * 1.5;      // power-of-two rounding of every partition's table
    const double target = std::max<double>((double)budget / 4.0, 1 << 20);
    int bits = 1;
    while (bits < RP_MAX_BITS && table_bytes / (double)(1 << bits) > target) bits++;
    return bits;
}

void radix_build(H *h, const void *keys, const uint8_t *validity, int kw, int64_t n, float lf, int64_t budget) {
    const int bits = choose_radix_bits(h, n, lf, budget);
    const int P = 1 << bits;
    h->radix_bits = bits;
    void *pk = nullptr, *prid = nullptr;
    std::vector<void *> temps;
    try {
        CU(cudaMallocAsync(&pk, (size_t)n * kw + 16, h->stream));
        temps.push_back(pk);
        CU(cudaMallocAsync(&prid, (size_t)n * 4 + 16, h->stream));
        temps.push_back(prid);
        radix_partition(h, keys, validity, 0, kw, n, bits, pk, (uint32_t *)prid, &h->st.ms_build_partition);
        const unsigned long long *off = h->rscratch.h_off;
        const int64_t n_valid = (int64_t)off[P];

        PhaseTimer tm(h, &h->st.ms_build_insert);
        std::vector<uint64_t> base(P);
        std::vector<uint32_t> mask(P);
        uint64_t total = 0;
        for (int p = 0; p < P; p++) {
            const uint64_t cnt = off[p + 1] - off[p];
            uint64_t cap = 16;
            while ((double)cap * lf < (double)cnt) cap <<= 1;
            REQUIRE(cap - 1 <= 0xffffffffULL, B200J_ERR_UNSUPPORTED, "partition table too large");
            base[p] = total;
            mask[p] = (uint32_t)(cap - 1);
            total += cap;
        }
        h->n_slots = total;
        dev_alloc(h, h->slots, (size_t)total * sizeof(Slot));
        CU(cudaMemsetAsync(h->slots.p, 0, (size_t)total * sizeof(Slot), h->stream));
        dev_alloc(h, h->part_base, (size_t)P * 8);
        dev_alloc(h, h->part_mask, (size_t)P * 4);
        CU(cudaMemcpyAsync(h->part_base.p, base.data(), (size_t)P * 8, cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->part_mask.p, mask.data(), (size_t)P * 4, cudaMemcpyHostToDevice, h->stream));
        CU(cudaStreamSynchronize(h->stream));      // base/mask are stack vectors

        BuildArgs a{};
        a.hk = h->hk;
        a.ctr = h->d_ctr;
        a.keys = pk;
        a.validity = nullptr;
        a.rowid = (const uint32_t *)prid;
        a.n = n_valid;
        a.mis = 0;
        a.tv = h->view();
        if (n_valid > 0) {
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
