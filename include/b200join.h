/*
 * b200join.h — C ABI of libb200join.so, the B200-native (sm_100a) parallel hash join that drops in
 * behind the DataFusion operator surface of jamesfer/datafusion-parallelism.
 *
 * The library replaces, for ONE join, everything below `ParallelHashJoin::execute`
 * (reference src/operator/parallel_hash_join.rs:140-167):
 *
 *   reference (Rust, CPU)                                               this ABI
 *   -----------------------------------------------------------------   -----------------------------
 *   ParallelHashJoinExecutor::new            parallel_hash_join_executor.rs:72-84      b200j_create
 *   BuildImplementation::build_side          operator/build_implementation.rs:50-112   b200j_build_push (per batch)
 *     Version10::process_input_batch         version10/build_implementation.rs:60-71     "
 *     JoinStateInstance::add                 version10/parallel_join_execution_state.rs:91-133
 *     JoinStateInstance::compact             version10/parallel_join_execution_state.rs:135-213   b200j_build_finish
 *   calculate_hash                           src/shared/shared.rs:11-16                b200j_hash_keys (and fused in build/probe)
 *   lookup_*_probe_batch                     operator/probe_lookup_implementation/*.rs b200j_probe_batch
 *     get_matching_indices[_with_probe]      src/shared/shared.rs:29-81                  "  (index pairs)
 *     equal_rows_arr                         src/shared/datafusion_private.rs:40-80      "  (fused: slots hold key values)
 *     append_right_indices / semi / anti     src/shared/datafusion_private.rs:85-185     "
 *     take_multiple_record_batch             src/shared/shared.rs:83-92                b200j_gather
 *   emit_(un)matched_build_records           full.rs:181-201, left_outer.rs:174-195,
 *                                            left_semi.rs:166-179, left_anti.rs:166-179 b200j_probe_finish
 *   apply_join_filter_to_indices             src/shared/datafusion_private.rs:295-328  stays with the caller: take the
 *                                            pairs of b200j_probe_batch(B200J_OUT_MATCHES_ONLY), filter them, hand them back
 *                                            through b200j_apply_pairs, gather with b200j_gather.
 *
 * Rules of the boundary
 *   - plain C: pointers, sizes, PODs. No C++/torch/arrow types. All structs start with `struct_size`
 *     so the ABI can grow.
 *   - every function returns a status (0 = ok, < 0 = error). The message of the last error on this
 *     handle is b200j_last_error(h); b200j_last_error(NULL) is the last error of a failed create/free call
 *     on this thread. Nothing throws or aborts across the boundary.
 *   - input buffers are BORROWED for the duration of the call only. Results are OWNED by the handle
 *     until b200j_result_release.
 *   - there is NO CPU fallback: without a CUDA device every compute call returns B200J_ERR_CUDA.
 *   - handle methods may be called concurrently from different threads for distinct `partition`
 *     ids; the library serialises them onto its CUDA streams.
 */
#ifndef B200JOIN_H
#define B200JOIN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200J_ABI_VERSION 1

/* status codes */
#define B200J_OK             0
#define B200J_ERR_INVALID   -1   /* bad argument / call order (e.g. probe before build finished)       */
#define B200J_ERR_CUDA      -2   /* CUDA runtime / driver error, or no device                          */
#define B200J_ERR_OOM       -3   /* device or pinned-host allocation failed                            */
#define B200J_ERR_UNSUPPORTED -4 /* valid request the library does not implement (yet)                 */
#define B200J_ERR_STATE     -5   /* "State already consumed" (version10/build_implementation.rs:32-33) */
#define B200J_ERR_NCCL      -6

/* DataFusion JoinType discriminants (datafusion-common 40, join_type.rs order) */
#define B200J_INNER       0
#define B200J_LEFT        1
#define B200J_RIGHT       2
#define B200J_FULL        3
#define B200J_LEFT_SEMI   4
#define B200J_RIGHT_SEMI  5
#define B200J_LEFT_ANTI   6
#define B200J_RIGHT_ANTI  7

/* where a batch's buffers live */
#define B200J_MEM_HOST    0      /* pageable or pinned host memory: staged to the device in chunks     */
#define B200J_MEM_DEVICE  1      /* device pointers on the handle's device (north_star: "as device pointers") */

/* column type: a fixed width in bytes (1,2,4,8,16), or B200J_T_UTF8 (32-bit offsets + data) */
#define B200J_T_UTF8      0

/* table layout chosen at build_finish */
#define B200J_LAYOUT_AUTO    0   /* direct while the table fits the L2 budget, else radix-partitioned  */
#define B200J_LAYOUT_DIRECT  1   /* one open-addressing table                                          */
#define B200J_LAYOUT_RADIX   2   /* radix-partition both sides, one table per partition (L2-resident)  */

/* NULL markers in index outputs (an Arrow validity bitmap is produced as well when requested) */
#define B200J_NULL_BUILD  UINT64_MAX
#define B200J_NULL_PROBE  UINT32_MAX

/* b200j_probe_batch `want` flags */
#define B200J_OUT_PAIRS         1u   /* (build_idx u64, probe_idx u32) index pairs — ProbeBuildIndices, shared.rs:24-27 */
#define B200J_OUT_HOST          2u   /* results in pinned host memory instead of device memory           */
#define B200J_OUT_MATCHES_ONLY  4u   /* only steps 1-3 (hash, lookup, equality): no join-type post-processing,
                                        no visited marking — the caller applies a JoinFilter and comes back
                                        through b200j_apply_pairs                                          */
#define B200J_OUT_COUNT_ONLY    8u   /* count the result rows, write nothing                              */
#define B200J_OUT_BUILD_U32     16u  /* build indices as uint32_t (a device holds < 2^31 build rows): b200j_result.build_idx then
                                        points at uint32_t values, NULL side = UINT32_MAX, build_idx_width = 4. A host result
                                        crosses PCIe with 8 instead of 12 bytes per pair; arrow's take accepts UInt32 indices */

#define B200J_OUT_GATHER_HINT   32u  /* the device result will be handed to b200j_gather: the library may keep auxiliary device
                                        arrays with it until b200j_result_release -- the partitioned probe keys (4 or 8 B per row)
                                        and, once the handle has seen two probe rows per build row, the table slot of every pair
                                        (4 B per row) -- and then keeps slot-ordered copies of the fixed-width build payload
                                        columns in the handle; the gather then reads neither the probe key column nor the build
                                        store at random. Ignored with B200J_OUT_HOST / COUNT_ONLY and
                                        wherever the dense radix probe does not apply; the pairs themselves are unchanged.
                                        A handle whose device pairs have been through b200j_gather once treats every later
                                        device probe as if it carried this flag (environment B200J_AUTO_HINT=0 turns that off)   */

#define B200J_OUT_PROBE_ORDER   64u  /* unique build keys (every probe row has at most one match -- a foreign-key join): the result is
                                        ONE array in the order of the probe batch, build_idx[i] = the build row probe row i matches
                                        or the NULL build index, and no probe_idx array (result.probe_order = 1, probe_idx = NULL,
                                        n = the batch's rows; Inner / Left callers drop the NULL rows, the probe columns need no
                                        take at all). 4 bytes per probe row back over PCIe with B200J_OUT_BUILD_U32 instead of 8.
                                        Honoured for Inner / Left / Right / Full joins on a single fixed-width key without
                                        B200J_OUT_MATCHES_ONLY / COUNT_ONLY when the build side has no duplicate keys; otherwise
                                        the call returns ordinary pairs (probe_order = 0)                                       */

typedef struct b200j_handle b200j_handle;

/* One Arrow array's buffers (Arrow columnar format: LSB-first validity bitmap, 32-bit Utf8 offsets). */
typedef struct {
    const void    *values;      /* fixed width: values buffer; Utf8: data buffer                     */
    const int32_t *offsets;     /* Utf8 only: n_rows + 1 offsets; NULL otherwise                     */
    const uint8_t *validity;    /* NULL => no nulls                                                  */
    int64_t        bit_offset;  /* offset (in rows) of row 0 inside `validity` (Arrow array offset)  */
} b200j_column;

typedef struct {
    uint32_t            struct_size;
    int32_t             mem;          /* B200J_MEM_*                                                  */
    int64_t             n_rows;
    int32_t             n_keys;       /* must equal the config's n_keys                              */
    int32_t             n_payload;    /* must equal the config's payload count of that side, or 0    */
    const b200j_column *keys;
    const b200j_column *payload;
} b200j_batch;

typedef struct {
    uint32_t  struct_size;
    int32_t   join_type;              /* B200J_INNER ...                                              */
    int32_t   n_partitions;           /* number of DataFusion partitions that will call build/probe   */
    int32_t   device;                 /* CUDA device ordinal                                          */
    int32_t   n_keys;
    const int32_t *key_widths;        /* [n_keys] 4 or 8 (Int32/UInt32/Date32 | Int64/UInt64/Timestamp ...) or B200J_T_UTF8
                                         (Utf8: b200j_column.offsets = 32-bit Arrow offsets, .values = data bytes).
                                         n_keys > 1 (composite key) or a Utf8 key: the table is keyed on the row hash
                                         (create_hashes, shared.rs:11-16) and key equality is checked on the candidate
                                         pairs, as the reference does (equal_rows_arr, datafusion_private.rs:40-80)      */
    int32_t   n_build_payload;
    const int32_t *build_payload_types;   /* [n_build_payload] width or B200J_T_UTF8                  */
    int32_t   n_probe_payload;
    const int32_t *probe_payload_types;
    uint64_t  hash_keys[4];           /* ahash RandomState keys; all zero => RandomState::with_seed(0) over the
                                         fixed PI seeds (shared.rs:14)                                */
    int32_t   hash_flags;             /* bit0: 0 = `specialize` integer path (nightly, as the reference), 1 = stable;
                                         bit1 (tests only): composite-key row hashes truncated to 10 bits, so that distinct
                                         keys collide and the equal_rows filter is exercised                          */
    int32_t   layout;                 /* B200J_LAYOUT_*                                               */
    int32_t   radix_bits;             /* 0 = choose; else log2(number of radix partitions)            */
    float     load_factor;            /* 0 = default 0.5                                              */
    int64_t   expected_build_rows;    /* sizing hint, 0 = unknown                                     */
    int64_t   l2_budget_bytes;        /* 0 = 0.9 x cudaDevAttrL2CacheSize (measured crossover) */
    void     *stream;                 /* cudaStream_t to run on; NULL = the library creates its own   */
    uint32_t  flags;                  /* B200J_FLAG_*                                                 */
} b200j_config;

#define B200J_FLAG_TIMING  1u         /* record CUDA-event timings of every phase into b200j_stats     */

/* Result of one probe call (or of the finaliser). Arrays are device or pinned-host memory according to
 * B200J_OUT_HOST. `n` rows; build_idx[i] == B200J_NULL_BUILD / probe_idx[i] == B200J_NULL_PROBE mark the
 * NULL side of outer rows; semi/anti results fill only the side they output. */
typedef struct {
    uint32_t  struct_size;
    int32_t   mem;
    int64_t   n;
    uint64_t *build_idx;
    uint32_t *probe_idx;
    int64_t   n_build_null;           /* rows whose build side is NULL                                */
    int64_t   n_probe_null;
    void     *_owner;                 /* private                                                      */
    int32_t   build_idx_width;        /* 8 (uint64_t, the default) or 4 (B200J_OUT_BUILD_U32)          */
    int32_t   probe_order;            /* 1: B200J_OUT_PROBE_ORDER was honoured -- probe_idx is NULL, n = the batch's rows and
                                         build_idx[i] is the match of probe row i (or the NULL build index)   */
} b200j_result;

/* Gathered output column (Arrow layout), produced by b200j_gather. */
typedef struct {
    void     *values;
    int32_t  *offsets;
    uint8_t  *validity;               /* NULL when the column has no nulls                            */
    int64_t   null_count;
    int64_t   values_bytes;
} b200j_out_column;

typedef struct {
    uint32_t          struct_size;
    int32_t           mem;
    int64_t           n_rows;
    int32_t           n_cols;         /* build key+payload columns first, then probe key+payload columns
                                         (build_join_schema, probe_lookup_implementation.rs:77-94)    */
    b200j_out_column *cols;
    void             *_owner;
} b200j_gathered;

typedef struct {
    uint32_t struct_size;
    int32_t  layout;                  /* layout actually used                                         */
    int64_t  build_rows;
    int64_t  build_null_keys;
    int64_t  distinct_keys;
    int64_t  duplicate_rows;          /* build rows whose key was already present                     */
    int64_t  table_slots;
    int64_t  table_bytes;
    int32_t  radix_bits;
    int32_t  _pad;
    int64_t  probe_rows;              /* accumulated over probe calls                                 */
    int64_t  result_rows;
    /* device time, ms (CUDA events on the library's stream), accumulated; valid with B200J_FLAG_TIMING */
    double   ms_build_copy;           /* concat of pushed batches (compact, a3)                       */
    double   ms_build_partition;      /* radix histogram + scatter of the build side                  */
    double   ms_build_insert;         /* table clear + hash + CAS insert (+ duplicate CSR)            */
    double   ms_probe_partition;      /* radix histogram + scatter of probe batches                   */
    double   ms_probe;                /* hash + lookup + pair compaction (the dominant kernel)        */
    double   ms_gather;
    int64_t  probe_launches;          /* launches of the dominant probe kernel                        */
    int64_t  kernel_launches;         /* all kernel launches of this handle                           */
    int64_t  probe_retries;           /* probe kernels re-run because the pair buffer was too small   */
    int64_t  dense_probe_launches;    /* probe launches that took the dense path (unique build keys behind the
                                         partition pass: 8 B key in, 8 B build index out per row, the partitioned
                                         row-id array is the result's probe-index column)               */
    int64_t  compact_launches;        /* dense results with misses compacted afterwards (Inner / Left)  */
    double   ms_exchange_partition;   /* b200j_shuffle_count + b200j_shuffle_scatter kernels             */
    int64_t  seg_probe_launches;      /* dense probe launches over the block-locally partitioned stream (probe_dense_seg_kernel) */
    int64_t  narrow_probe_launches;   /* ... of which the stream carried 8-byte keys as 32-bit values                    */
    int64_t  gather_seg_launches;     /* b200j_gather launches that walked the build side partition by partition (gather_seg_multi_kernel) */
    int64_t  gather_hint_launches;    /* ... and launches that used what B200J_OUT_GATHER_HINT kept (stream keys, slot-ordered payload)    */
    double   ms_slot_order;           /* building the slot-ordered payload copies (once per build, on the first hinted gather)            */
    int64_t  probe_order_results;     /* probe calls answered in probe order (B200J_OUT_PROBE_ORDER)                                      */
} b200j_stats;

/* ---- lifecycle ------------------------------------------------------------------------------- */
int32_t     b200j_abi_version(void);
int32_t     b200j_device_count(void);                       /* 0 when no CUDA device is visible    */
int32_t     b200j_create(const b200j_config *cfg, b200j_handle **out);
int32_t     b200j_destroy(b200j_handle *h);
const char *b200j_last_error(const b200j_handle *h);
int32_t     b200j_get_stats(b200j_handle *h, b200j_stats *out);
int32_t     b200j_reset_stats(b200j_handle *h);
int32_t     b200j_synchronize(b200j_handle *h);             /* wait for the handle's stream        */
/* Re-arm the handle for another execution of the same join (the plan being executed again,
 * ParallelHashJoin::execute on a fresh OnceLock, parallel_hash_join.rs:144-151): drops the build
 * side and the table, keeps streams and device allocations. Outstanding results stay valid. */
int32_t     b200j_reset(b200j_handle *h);
/* The handle keeps freed device blocks of 1 MiB and more for reuse (a cudaMallocAsync of an 8 GB block costs as much
 * as the probe kernel), up to 40 % of the device's memory (B200J_CACHE_LIMIT_BYTES overrides). b200j_trim gives every
 * cached free block back to the driver: call it when another allocator in the process needs the room. */
int32_t     b200j_trim(b200j_handle *h);

/* ---- build side ------------------------------------------------------------------------------ */
/* Append one build batch (rows are numbered in arrival order, like offset_counter.fetch_add,
 * parallel_join_execution_state.rs:95). */
int32_t b200j_build_push(b200j_handle *h, int32_t partition, const b200j_batch *batch);
/* End of this partition's build stream. The LAST partition to call runs the build
 * (partition → insert → duplicate CSR); the others return immediately.  A partition calling
 * twice gets B200J_ERR_STATE. */
int32_t b200j_build_finish(b200j_handle *h, int32_t partition);

/* ---- probe side ------------------------------------------------------------------------------ */
int32_t b200j_probe_batch(b200j_handle *h, int32_t partition, const b200j_batch *batch, uint32_t want,
                          b200j_result *out);
/* One probe batch handed over as several device-resident pieces of its (single, fixed-width, non-NULL) key column: the
 * batch is their concatenation, probe_idx counts across the pieces. A piece may live in a PEER GPU's memory mapped with
 * b200j_ipc_open: the radix-partition pass of the local join then pulls it over NVLink while it partitions, which is how
 * the multi-GPU exchange is fused into the join (DistributedJoin mode "pull": the pieces are the regions the source ranks
 * scattered for this rank with b200j_shuffle_local_scatter). */
int32_t b200j_probe_pieces(b200j_handle *h, int32_t partition, int32_t n_pieces, const void *const *piece_keys, const int64_t *piece_rows,
                           int32_t piece_key_width /* 0 = the join's key width; 4 with 8-byte join keys = zero-extended 32-bit values */,
                           uint32_t want, b200j_result *out);
/* Join-type post-processing (a7) of caller-filtered pairs of one probe batch of `n_probe_rows` rows. */
int32_t b200j_apply_pairs(b200j_handle *h, int32_t partition, int64_t n_probe_rows, int32_t mem, int64_t n_pairs,
                          const uint64_t *build_idx, const uint32_t *probe_idx, uint32_t want, b200j_result *out);
/* End of this partition's probe stream. The LAST partition gets the finaliser rows of
 * Left / Full / LeftSemi / LeftAnti (build rows, probe side NULL); everybody else gets n = 0. */
int32_t b200j_probe_finish(b200j_handle *h, int32_t partition, uint32_t want, b200j_result *out);
int32_t b200j_result_release(b200j_handle *h, b200j_result *r);

/* Output gather (a6): build columns at build_idx, then probe columns of `probe_batch` at probe_idx.
 * probe_batch may be NULL for the finaliser (probe columns all NULL). A B200J_OUT_PROBE_ORDER result is gathered with
 * probe_batch = NULL: its build columns come out in the order of the probe batch, whose own columns are the output's
 * probe columns as they are. */
int32_t b200j_gather(b200j_handle *h, const b200j_result *pairs, const b200j_batch *probe_batch, uint32_t want,
                     b200j_gathered *out);
int32_t b200j_gathered_release(b200j_handle *h, b200j_gathered *g);

/* ---- coalescing feeder for small probe batches (replaces what WorkStealingRepartitionExec,
 * work_stealing_repartition_exec.rs:204-295, does in front of the reference's operator) ------------------------------
 * DataFusion hands over 8192-row batches; a GPU probe wants millions of rows per launch. b200j_probe_push COPIES one
 * batch (key and probe payload columns, host or device) to the end of the partition's stage on the device; rows are
 * numbered consecutively since the partition's last flush. b200j_probe_flush probes everything staged in one go
 * (probe_idx = staged row number); b200j_gather_staged gathers with the staged probe columns -- the caller slices the
 * gathered columns into output RecordBatches of whatever size it likes -- and empties the stage. Key columns must be
 * fixed-width (Utf8 payload is fine). b200j_staged_rows tells the caller when it is worth flushing. */
int32_t b200j_probe_push(b200j_handle *h, int32_t partition, const b200j_batch *batch);
int64_t b200j_staged_rows(b200j_handle *h, int32_t partition);
int32_t b200j_probe_flush(b200j_handle *h, int32_t partition, uint32_t want, b200j_result *out);
int32_t b200j_gather_staged(b200j_handle *h, int32_t partition, const b200j_result *pairs, uint32_t want, b200j_gathered *out);

/* ---- multi-GPU exchange (no counterpart in the reference, which is single-process) ----------
 * One process per GPU; every rank hash-partitions BOTH join sides by destination rank
 *     dest = (((hash >> 32) & 0xfffff) * n_dest) >> 20        (independent of the radix and slot bits)
 * and the scatter kernel writes each (key, row id) straight into the destination's receive buffer: the
 * per-destination base pointers may be peer-GPU memory mapped over NVLink (b200j_ipc_*), in which case the
 * scatter IS the all-to-all, or regions of a local send buffer handed to ncclSend/Recv by the caller.
 * Call order per batch: b200j_shuffle_count -> (ranks exchange the counts) -> b200j_shuffle_scatter.
 * No other call on the handle may come between the two. Rows whose key is NULL are counted in
 * counts[n_dest] and written to destination n_dest (keep that one local: NULL keys never match). */
#define B200J_MAX_DEST 16
int32_t b200j_shuffle_count(b200j_handle *h, const b200j_batch *batch, int32_t n_dest, int64_t *counts /* [n_dest + 1], [n_dest + 2] with a hot set */);
int32_t b200j_shuffle_scatter(b200j_handle *h, int32_t n_dest, void *const *dest_keys /* [n_dest + 1] */,
                              uint32_t *const *dest_rowid /* [n_dest + 1] */, const int64_t *dest_offset /* [n_dest + 1] first row to
                              write in each destination buffer */, uint32_t rowid_base /* added to the batch-local row index */);
/* Single-pass variant for exchanges carried by the copy engines (round 2): ONE kernel reads the batch once and writes every
 * (key, row id) into the region of its destination inside LOCAL send buffers -- send_keys / send_rowid hold n_dest + 1 (+ 1
 * with a hot set) regions of `region_rows` rows each, in the order [ranks][NULL keys][hot keys]. No histogram pass: a tile
 * reserves its run in a region with one atomic. counts[] receives the rows written per region. The regions are then handed
 * to the peers with b200j_copy_async (cudaMemcpyAsync over NVLink: no SM is busy while the local join of the previous
 * chunk runs). Returns B200J_ERR_INVALID with "region too small" when a region overflowed (nothing past a region's end is
 * written): call again with larger regions. */
int32_t b200j_shuffle_local_scatter(b200j_handle *h, const b200j_batch *batch, int32_t n_dest, void *send_keys, uint32_t *send_rowid,
                                    int64_t region_rows, uint32_t rowid_base, int32_t key_out_width,
                                    int64_t *counts /* [n_dest + 1], [n_dest + 2] with a hot set */);
/* key_out_width: 0 / the key width = keys as they are; 4 with 8-byte keys = the regions take the keys as 32-bit values (half
 * the bytes the peers pull over NVLink; b200j_probe_pieces takes such pieces with piece_key_width = 4). A key outside
 * [0, 2^32) then fails the call with B200J_ERR_INVALID "keys do not fit 32 bits": repeat it with key_out_width = 8. */
/* Enqueue a device-to-device copy (same device, or a peer mapped with b200j_ipc_open) on one of the handle's copy lanes
 * (0..3); it starts after the work enqueued on the handle's stream so far. b200j_copy_wait blocks until every lane is idle. */
int32_t b200j_copy_async(b200j_handle *h, void *dst, const void *src, size_t bytes, int32_t lane);
int32_t b200j_copy_wait(b200j_handle *h);
/* Heavy-hitter routing (SURVEY §8e "skew"). Rows whose key is in the hot set are NOT sent to the key's owner: they go
 * to one extra, rank-local destination, to be joined there against the build rows of the hot keys, which the caller
 * broadcasts (a hot key would otherwise deliver its whole probe volume to one GPU). With a non-empty hot set the count /
 * scatter calls take n_dest + 2 entries: [0, n_dest) ranks, [n_dest] NULL keys, [n_dest + 1] hot keys. `keys`: host array
 * of key values (32-bit keys zero-extended); the value 0 is ignored; n = 0 switches the routing off. */
int32_t b200j_shuffle_set_hot(b200j_handle *h, const uint64_t *keys, int32_t n);
/* Device buffers that peers on the same node can map (cudaIpc): `handle_out` is 64 opaque bytes to send to the peers. */
#define B200J_IPC_HANDLE_BYTES 64
int32_t b200j_ipc_alloc(b200j_handle *h, size_t bytes, void **ptr, uint8_t *handle_out);
int32_t b200j_ipc_open(b200j_handle *h, const uint8_t *handle, void **peer_ptr);
int32_t b200j_ipc_close(b200j_handle *h, void *peer_ptr);
int32_t b200j_ipc_free(b200j_handle *h, void *ptr);

/* ---- output rows across GPUs (a6 over the exchange; the reference is single-process) ---------------------------
 * take_multiple_record_batch (src/shared/shared.rs:83-92) when the join's sides were handed over as PIECES -- one piece per source
 * rank of the multi-GPU exchange: the local join's indices count the rows this rank RECEIVED (build rows in build_push order,
 * probe rows in piece order), while the columns stayed on the rank that owns the row, mapped into this process with
 * b200j_ipc_open (or plain device pointers when several "ranks" share a GPU). Late materialisation in two steps, every array in
 * DEVICE memory, every call enqueued on the handle's stream:
 *   b200j_resolve_origin   index -> (piece, row at the piece's source). idx: u32 / u64 indices (idx_width 4 / 8) into the
 *                          concatenation of the pieces; a NULL index (all ones) or one past the pieces gives row 0xffffffff.
 *                          piece_rowid[p]: the u32 row ids that travelled with piece p (b200j_shuffle_* wrote them; may be peer
 *                          memory; NULL = the position inside the piece), piece_rowid_base[p] is added (NULL = zeros).
 *   b200j_gather_pieces    one fixed-width column (1/2/4/8/16 bytes): out_values[i] = col_values[piece[i]][row[i]], validity
 *                          word i/32 bit i%32 = row is not 0xffffffff and the source row is valid (col_validity[p]: LSB bitmap
 *                          from the column's row 0, or NULL; the array itself may be NULL). out_validity may be NULL.
 *   b200j_gather_pieces_utf8_offsets / _data   one Utf8 column (int32 Arrow offsets + bytes per source): the first call writes the
 *                          n + 1 output offsets and the validity words and RETURNS the byte total (it synchronises the stream);
 *                          the caller allocates that many bytes and the second call copies them.
 *   b200j_utf8_rows_equal  out[i] = 1 when row i of two gathered Utf8 columns holds the same bytes and both are valid: the key
 *                          check of a join keyed on row hashes (equal_rows_arr, src/shared/datafusion_private.rs:40-80).
 * At most B200J_MAX_PIECES pieces per call. */
#define B200J_MAX_PIECES 64
int32_t b200j_resolve_origin(b200j_handle *h, int64_t n, const void *idx, int32_t idx_width, int32_t n_pieces, const int64_t *piece_rows,
                             const uint32_t *const *piece_rowid, const uint32_t *piece_rowid_base, uint32_t *out_piece, uint32_t *out_row);
int32_t b200j_gather_pieces(b200j_handle *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces, const void *const *col_values,
                            const uint8_t *const *col_validity, int32_t width, void *out_values, uint32_t *out_validity);
int32_t b200j_gather_pieces_utf8_offsets(b200j_handle *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces,
                                         const int32_t *const *col_offsets, const uint8_t *const *col_validity, int32_t *out_offsets,
                                         uint32_t *out_validity, int64_t *total_bytes);
int32_t b200j_gather_pieces_utf8_data(b200j_handle *h, int64_t n, const uint32_t *piece, const uint32_t *row, int32_t n_pieces,
                                      const int32_t *const *col_offsets, const uint8_t *const *col_data, const int32_t *out_offsets, uint8_t *out_data);
int32_t b200j_utf8_rows_equal(b200j_handle *h, int64_t n, const int32_t *off_a, const uint8_t *data_a, const uint32_t *valid_a, const int32_t *off_b,
                              const uint8_t *data_b, const uint32_t *valid_b, uint8_t *out);

/* ---- allocation-free destination count -------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same result as b200j_shuffle_count
 * (rows of one 4/8-byte key column per destination rank, counts[n_dest] = NULL-key rows), but
 *   - the call allocates NOTHING, on the device or through any allocating library routine: the per-CTA histograms the count
 *     kernel needs live in a file-scope __device__ array of the library (272 KiB, held by every GPU that loads the library
 *     until the process exits), whose device address is looked up once per handle in b200j_create;
 *   - inputs and outputs are the CALLER's device buffers: `keys` (n_rows values of the handle's key width, aligned to that
 *     width), `validity` (Arrow LSB bitmap with `bit_offset`, or NULL), `counts` (n_dest + 1 int64 values, DEVICE memory);
 *   - the work is only ENQUEUED on the handle's stream (no synchronisation, no host copy): read `counts` after
 *     b200j_synchronize or after a stream-ordered copy of your own;
 *   - at most B200J_STATIC_MAX_ROWS rows per call: more fails with B200J_ERR_INVALID "n_rows exceeds
 *     B200J_STATIC_MAX_ROWS" (split the batch and add the counts);
 *   - a hot-key set (b200j_shuffle_set_hot) is not supported here: B200J_ERR_STATE while one is set; it does not arm
 *     b200j_shuffle_scatter (that call still wants a b200j_shuffle_count).
 * EVERY call on EVERY handle of a device shares the one array, and it keeps the last call's contents. Nothing in it has to
 * start at zero (each call overwrites every entry it later reads), but two calls whose kernels overlap on the GPU -- two
 * handles with different streams, or one stream forked by the caller -- race on it: keep such calls ordered. */
#define B200J_STATIC_MAX_ROWS 134217728          /* 2^27 = 4096 CTAs x 32768 rows: what the static array has room for */
int32_t b200j_shuffle_count_static(b200j_handle *h, const void *keys, const uint8_t *validity, int64_t bit_offset, int64_t n_rows,
                                   int32_t n_dest, int64_t *counts /* device, [n_dest + 1] */);

/* ---- hashing (a1) ---------------------------------------------------------------------------- */
/* create_hashes over the key columns of `batch` with the handle's RandomState: one u64 per row into
 * `out` (same memory space as the batch). Bit-identical to oracle/join_oracle.c:orc_create_hashes. */
int32_t b200j_hash_keys(b200j_handle *h, const b200j_batch *batch, uint64_t *out);

/* ---- pinned host memory for callers that want full-speed PCIe copies -------------------------- */
int32_t b200j_host_alloc(size_t bytes, void **out);
int32_t b200j_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif /* B200JOIN_H */
