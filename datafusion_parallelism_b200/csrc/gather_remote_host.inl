// gather_remote_host.inl — host side of gather_remote.cuh (included inside b200join.cu's anonymous namespace).

void do_resolve_origin(H *h, int64_t n, const void *idx, int32_t idx_width, int32_t n_pieces, const int64_t *piece_rows,
                       const uint32_t *const *piece_rowid, const uint32_t *piece_rowid_base, uint32_t *out_piece, uint32_t *out_row) {
    REQUIRE(n >= 0, B200J_ERR_INVALID, "n is negative");
    REQUIRE(idx_width == 4 || idx_width == 8, B200J_ERR_INVALID, "idx_width must be 4 or 8");
    REQUIRE(n_pieces >= 1 && n_pieces <= GP_MAX_PIECES, B200J_ERR_INVALID, "n_pieces must be in 1..64");
    REQUIRE(piece_rows != nullptr, B200J_ERR_INVALID, "piece_rows is NULL");
    REQUIRE(n == 0 || (idx && out_piece && out_row), B200J_ERR_INVALID, "idx / out_piece / out_row is NULL");
    if (n == 0) return;
    OriginPieces op{};
    op.n = n_pieces;
    int64_t end = 0;
    for (int p = 0; p < n_pieces; p++) {
        REQUIRE(piece_rows[p] >= 0, B200J_ERR_INVALID, "negative piece_rows");
        end += piece_rows[p];
        REQUIRE(end < 0xffffffffLL, B200J_ERR_UNSUPPORTED, "the pieces hold more than 2^32 - 1 rows");
        op.end[p] = (uint32_t)end;
        op.rowid[p] = piece_rowid ? piece_rowid[p] : nullptr;
        op.rowid_base[p] = piece_rowid_base ? piece_rowid_base[p] : 0u;
    }
    const int grid = grid_for(h, (n + 127) / 128, 8, 8);
    if (idx_width == 8)
        resolve_origin_kernel<unsigned long long><<<grid, 256, 0, h->stream>>>(op, (const unsigned long long *)idx, n, (uint32_t)end, out_piece, out_row);
    else
        resolve_origin_kernel<uint32_t><<<grid, 256, 0, h->stream>>>(op, (const uint32_t *)idx, n, (uint32_t)end, out_piece, out_row);
    h->st.kernel_launches++;
    CU(cudaGetLastError());
}

PieceCols piece_cols(int32_t n_pieces, const void *const *values, const uint8_t *const *validity, const uint8_t *const *data) {
    REQUIRE(n_pieces >= 1 && n_pieces <= GP_MAX_PIECES, B200J_ERR_INVALID, "n_pieces must be in 1..64");
    REQUIRE(values != nullptr, B200J_ERR_INVALID, "the column pointer array is NULL");
    PieceCols pc{};
    pc.n = n_pieces;
    for (int p = 0; p < n_pieces; p++) {
        pc.values[p] = values[p];
        pc.validity[p] = validity ? validity[p] : nullptr;
        pc.data[p] = data ? data[p] : nullptr;
    }
    return pc;
}

void do_gather_pieces(H *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces, const void *const *col_values,
                      const uint8_t *const *col_validity, int32_t width, void *out_values, uint32_t *out_validity) {
    REQUIRE(n >= 0, B200J_ERR_INVALID, "n is negative");
    REQUIRE(width == 1 || width == 2 || width == 4 || width == 8 || width == 16, B200J_ERR_UNSUPPORTED, "width must be 1, 2, 4, 8 or 16");
    const PieceCols pc = piece_cols(n_pieces, col_values, col_validity, nullptr);
    REQUIRE(n == 0 || (piece && row && out_values), B200J_ERR_INVALID, "piece / row / out_values is NULL");
    if (n == 0) return;
    const int grid = grid_for(h, (n + 127) / 128, 8, 8);
    switch (width) {
        case 1: gather_pieces_kernel<uint8_t><<<grid, 256, 0, h->stream>>>(pc, piece, row, n, (uint8_t *)out_values, out_validity, nullptr); break;
        case 2: gather_pieces_kernel<uint16_t><<<grid, 256, 0, h->stream>>>(pc, piece, row, n, (uint16_t *)out_values, out_validity, nullptr); break;
        case 4: gather_pieces_kernel<uint32_t><<<grid, 256, 0, h->stream>>>(pc, piece, row, n, (uint32_t *)out_values, out_validity, nullptr); break;
        case 8: gather_pieces_kernel<unsigned long long><<<grid, 256, 0, h->stream>>>(pc, piece, row, n, (unsigned long long *)out_values, out_validity, nullptr); break;
        default: gather_pieces_kernel<ulonglong2><<<grid, 256, 0, h->stream>>>(pc, piece, row, n, (ulonglong2 *)out_values, out_validity, nullptr); break;
    }
    h->st.kernel_launches++;
    CU(cudaGetLastError());
}

void do_gather_pieces_utf8_offsets(H *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces, const int32_t *const *col_offsets,
                                   const uint8_t *const *col_validity, int32_t *out_offsets, uint32_t *out_validity, int64_t *total_bytes) {
    REQUIRE(n >= 0, B200J_ERR_INVALID, "n is negative");
    REQUIRE(out_offsets != nullptr && total_bytes != nullptr, B200J_ERR_INVALID, "out_offsets / total_bytes is NULL");
    const PieceCols pc = piece_cols(n_pieces, (const void *const *)col_offsets, col_validity, nullptr);
    *total_bytes = 0;
    if (n == 0) {
        CU(cudaMemsetAsync(out_offsets, 0, 4, h->stream));
        return;
    }
    REQUIRE(piece && row, B200J_ERR_INVALID, "piece / row is NULL");
    std::vector<void *> temps;
    try {
        int64_t *len = nullptr, *start = nullptr, *bsum = nullptr;
        const int64_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
        CU(big_alloc(h, (void **)&len, (size_t)n * 8)); temps.push_back(len);
        CU(big_alloc(h, (void **)&start, (size_t)n * 8)); temps.push_back(start);
        CU(big_alloc(h, (void **)&bsum, (size_t)(n_tiles + 1) * 8)); temps.push_back(bsum);
        int64_t *dtotal = bsum + n_tiles;
        gather_pieces_utf8_len_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>(pc, piece, row, n, len, out_validity, nullptr);
        scan_block_sums_kernel<<<(unsigned)n_tiles, SCAN_THREADS, 0, h->stream>>>(len, n, bsum);
        scan_serial_kernel<<<1, 1024, 0, h->stream>>>(bsum, n_tiles, dtotal);
        scan_apply_kernel<<<(unsigned)n_tiles, SCAN_THREADS, 0, h->stream>>>(len, n, bsum, start, out_offsets);
        h->st.kernel_launches += 4;
        CU(cudaGetLastError());
        int64_t total = 0;
        CU(cudaMemcpyAsync(&total, dtotal, 8, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        REQUIRE(total <= 0x7fffffffLL, B200J_ERR_UNSUPPORTED, "gathered Utf8 column exceeds 2 GiB (32-bit Arrow offsets): gather fewer rows per call");
        const int32_t t32 = (int32_t)total;
        CU(cudaMemcpyAsync(out_offsets + n, &t32, 4, cudaMemcpyHostToDevice, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        *total_bytes = total;
        free_temps(h, temps);
    } catch (...) {
        free_temps(h, temps);
        throw;
    }
}

void do_gather_pieces_utf8_data(H *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces, const int32_t *const *col_offsets,
                                const uint8_t *const *col_data, const int32_t *out_offsets, uint8_t *out_data) {
    REQUIRE(n >= 0, B200J_ERR_INVALID, "n is negative");
    const PieceCols pc = piece_cols(n_pieces, (const void *const *)col_offsets, nullptr, col_data);
    if (n == 0) return;
    REQUIRE(piece && row && out_offsets && out_data && col_data, B200J_ERR_INVALID, "piece / row / out_offsets / out_data / col_data is NULL");
    gather_pieces_utf8_copy_kernel<<<grid_for(h, n * 32, 256, 16), 256, 0, h->stream>>>(pc, piece, row, n, out_offsets, out_data);
    h->st.kernel_launches++;
    CU(cudaGetLastError());
}

void do_utf8_rows_equal(H *h, int64_t n, const int32_t *off_a, const uint8_t *data_a, const uint32_t *valid_a, const int32_t *off_b,
                        const uint8_t *data_b, const uint32_t *valid_b, uint8_t *out) {
    REQUIRE(n >= 0, B200J_ERR_INVALID, "n is negative");
    if (n == 0) return;
    REQUIRE(off_a && off_b && out, B200J_ERR_INVALID, "offsets / out is NULL");
    utf8_rows_equal_kernel<<<grid_for(h, n, 256, 8), 256, 0, h->stream>>>(off_a, data_a, valid_a, off_b, data_b, valid_b, n, out);
    h->st.kernel_launches++;
    CU(cudaGetLastError());
}
