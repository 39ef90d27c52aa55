/*
 * drm_b200.h — C ABI of libdrm_cuda.so, the B200 (sm_100a) implementation of DeepReadMapper's
 * search-time hot path: batched HNSW-PQ search + Smith-Waterman rerank.
 *
 * The reference has no FFI layer; its hot path sits behind three C++ call sites.  Each entry point
 * below names the reference interface it replaces (paths relative to the reference repository):
 *
 *   drm_index_open / drm_index_close   faiss::read_index + dynamic_cast<IndexHNSWPQ*> + delete
 *                                      src/main.cpp:236-237,299 ; src/hnswpq/test_search.cpp:68-76
 *   drm_search                         faiss_search -> IndexHNSWPQ::search
 *                                      src/hnswpq/search.cpp:6-57 ; includes/hnswpq/search.hpp:18-22
 *   drm_ref_open_*                     ref_seqs / ref_genome arguments of post_process_sw_{static,dynamic}
 *                                      src/main.cpp:178-199 ; src/utils/post_processor.cpp:47-71
 *   drm_sw_rerank                      post_process_sw_static / post_process_sw_dynamic (+ find_sequences,
 *                                      sw_reranker)  src/utils/post_processor.cpp:357-549 ;
 *                                      src/utils/reranker.cpp:3-51
 *   drm_sw_pairs                       calc_sw_score over a batch of pairs (the test_sw_score workload)
 *                                      src/utils/metrics.cpp:10-45 ; src/test_sw_score.cpp:44-52
 *   drm_merge_topk_device              (new) per-read K-way merge for a position-sharded index
 *
 * Conventions: every function returns 0 on success and a negative DRM_E* code on failure;
 * drm_last_error() returns the thread-local message for the last failure on the calling thread.
 * The caller owns every host buffer; handles own device memory.  No C++ types, exceptions or torch
 * types cross this boundary.  There is NO CPU fallback: without a CUDA device every compute entry
 * point fails with DRM_ECUDA.
 *
 * "_device" variants take DEVICE pointers (already resident in HBM on the handle's device) and a
 * cudaStream_t passed as void*; they enqueue work and return without synchronising.
 */
#ifndef DRM_B200_H
#define DRM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DRM_OK 0
#define DRM_EINVAL (-1)      /* bad argument */
#define DRM_EIO (-2)         /* file missing / unreadable / not an IHNp index */
#define DRM_ECUDA (-3)       /* CUDA runtime error, or no device */
#define DRM_EUNSUPPORTED (-4) /* index parameters outside what the kernels implement */
#define DRM_ENOTENOUGH (-5)  /* "Not enough candidates (a < k)" — src/utils/reranker.cpp:26-29 */
#define DRM_EKTOOLARGE (-6)  /* "Final k too large..." — src/utils/post_processor.cpp:389-392,486-489 */
#define DRM_EEMPTY (-7)      /* "Query data is empty" — src/hnswpq/search.cpp:16-19 */

typedef struct drm_index drm_index;
typedef struct drm_ref drm_ref;
typedef struct drm_reads drm_reads;

/* Counters summed over the queries of one call (same definitions as FAISS hnsw_stats). */
typedef struct drm_search_stats {
    int64_t nhops0;    /* level-0 expansions */
    int64_t nhops_up;  /* upper-level row scans */
    int64_t ndis_eval; /* level-0 ADC evaluations performed by the kernel (>= FAISS ndis: the kernel
                          has no visited table) */
    double device_ms;  /* kernel time, CUDA events (host-pointer calls only; 0 for _device calls) */
    double h2d_ms;     /* host->device copy time (host-pointer calls only) */
    double d2h_ms;     /* device->host copy time (host-pointer calls only) */
} drm_search_stats;

const char *drm_last_error(void);
int drm_device_count(void);
const char *drm_version(void);

/* ------------------------------------------------------------------ index (FAISS IHNp file) */
/* devices: CUDA ordinals to load a full replica on (n_dev >= 1).  Host-pointer search calls shard
 * the query batch contiguously over the replicas; _device calls address one replica by slot. */
int drm_index_open(const char *index_file, const int *devices, int n_dev, drm_index **out);
void drm_index_close(drm_index *ix);
/* info[0..11] = d, ntotal, M_pq, nbits, M_hnsw, entry_point, max_level, efConstruction, efSearch,
 *               n_links, n_replicas, device bytes per replica */
int drm_index_info(const drm_index *ix, int64_t *info);

/* q: n x d row-major fp32 (host).  D: n x k fp32 ascending ADC squared-L2; I: n x k int64 labels,
 * unfilled slots D = FLT_MAX, I = -1.  stats may be NULL. */
int drm_search(drm_index *ix, const float *q, int64_t n, int d, int k, int ef, float *D, int64_t *I,
               drm_search_stats *stats);
/* Same on device pointers of replica `slot`; counters (3 x int64: nhops0, nhops_up, ndis_eval) are
 * accumulated into counters_dev when non-NULL. */
int drm_search_device(drm_index *ix, int slot, const float *q_dev, int64_t n, int d, int k, int ef,
                      float *D_dev, int64_t *I_dev, int64_t *counters_dev, void *stream);

/* ------------------------------------------------------------------ reference windows */
/* Genome form: window id -> (pos = id/2, reverse-complement = id%2), src/utils/post_processor.cpp:47-66.
 * genome bytes must be the cleaned sequence the reference holds (extract_FASTA_sequence). */
int drm_ref_open_genome(const char *genome, int64_t len, int ref_len, const int *devices, int n_dev,
                        drm_ref **out);
/* Table form: n_windows strings of ref_len bytes each (ref_seqs[id], find_sequence_static). */
int drm_ref_open_windows(const char *windows_flat, int64_t n_windows, int ref_len, const int *devices,
                         int n_dev, drm_ref **out);
void drm_ref_close(drm_ref *r);

/* ------------------------------------------------------------------ Smith-Waterman */
/* scores[i] = calc_sw_score(a_i, b_i); a_i = a_flat[a_off[i] .. a_off[i+1]). */
int drm_sw_pairs(const char *a_flat, const int64_t *a_off, const char *b_flat, const int64_t *b_off,
                 int64_t n, int32_t *scores);

/* The same scores for callers that may not allocate (code under CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  All pointers are DEVICE
 * pointers on `device` (offsets are int64, n + 1 of them); the launch is asynchronous on `stream`.  The call performs no
 * cudaMalloc / cudaFree / pool allocation: the kernel's DP rows live in a static device array of
 * DRM_SW_PAIRS_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit.  max_len_b = an upper bound on
 * the length of every b_i, at most DRM_SW_PAIRS_NOALLOC_MAX_LEN_B (DRM_EUNSUPPORTED above it); a pair whose b_i is longer
 * than max_len_b gets score -1.  Every call on a device shares the one array: calls that overlap on the GPU (two
 * streams, two host threads) race - serialise them.  Nothing in the array has to be cleared between calls. */
#define DRM_SW_PAIRS_NOALLOC_MAX_LEN_B 4095
#define DRM_SW_PAIRS_NOALLOC_SCRATCH_BYTES (16u << 20)
int drm_sw_pairs_device_noalloc(int device, const char *a_flat_dev, const int64_t *a_off_dev, const char *b_flat_dev,
                                const int64_t *b_off_dev, int64_t n, int max_len_b, int32_t *scores_dev, void *stream);

/* mode 0 = static bound rule (bound = number of dense windows), 1 = dynamic (bound = genome length;
 * stride == 1 keeps every id, windows that overrun the genome score 0).
 * neighbors: n x k_in int64 search labels; the first min(k_clusters, k_in) of each row are used.
 * Outputs: out_scores n x k (int32, descending, libstdc++ partial_sort tie order), out_ids n x k
 * (dense window ids, uint64), n_cands[n] (nullable).  Errors: DRM_EKTOOLARGE, DRM_ENOTENOUGH. */
int drm_sw_rerank(drm_ref *r, int mode, const int64_t *neighbors, int64_t n, int64_t k_in,
                  const char *q_flat, const int64_t *q_off, int64_t stride, int64_t k, int64_t k_clusters,
                  int32_t *out_scores, uint64_t *out_ids, int32_t *n_cands, double *device_ms);
/* Device form (replica `slot`): q_flat_dev/q_off_dev/neighbors_dev/out_* all device pointers;
 * max_qlen = max query length (host knows it).  min_cands_dev (int32, nullable) receives the minimum
 * candidate count over the batch so the caller can raise DRM_ENOTENOUGH after synchronising.
 * cells_dev (int64, nullable) accumulates the DP cell updates performed. */
int drm_sw_rerank_device(drm_ref *r, int slot, int mode, const int64_t *neighbors_dev, int64_t n, int64_t k_in,
                         const char *q_flat_dev, const int64_t *q_off_dev, int max_qlen, int64_t stride,
                         int64_t k, int64_t k_clusters, int32_t *out_scores_dev, uint64_t *out_ids_dev,
                         int32_t *n_cands_dev, int32_t *min_cands_dev, int64_t *cells_dev, void *stream);

/* ------------------------------------------------------------------ the whole stage in one call */
/* Search + SW rerank as the pipeline runs them back to back (src/main.cpp:270-296 then :301-354 with the
 * SW variants of :333-341): labels never return to the host between the two stages.  Reads are cut into
 * chunks that are double-buffered over two streams per replica, so host<->device copies overlap the
 * kernels (pass pinned host buffers for that to be effective).  The search asks for k_search labels
 * (K when stride == 1, K_clusters otherwise, src/main.cpp:55-63,278); the rerank uses the first
 * min(k_clusters, k_search) of them.  D/I (n x k_search) are the raw search output that the reference
 * saves as distances.npy / indices.npy; sw_scores/sw_ids (n x k) are the rerank output.  Any of the four
 * output pointers may be NULL to skip that copy.  timing_ms[0..1] (nullable) = {device span, wall}. */
int drm_search_rerank(drm_index *ix, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                      const char *q_flat, const int64_t *q_off, int64_t stride, int64_t k, int64_t k_clusters,
                      float *D, int64_t *I, int32_t *sw_scores, uint64_t *sw_ids, drm_search_stats *stats,
                      double *timing_ms);

/* ------------------------------------------------------------------ SW traceback (real CIGAR) */
/* The alignment behind a rerank score - what the reference's write_sam leaves as "TODO: Implement real CIGAR - from SW
 * ranker" (src/utils/utils.cpp:381-384).  ids: n x per_read window ids (e.g. the first columns of sw_ids); read i is
 * q_flat[q_off[i] .. q_off[i+1]) with `strip` bytes dropped from each end when longer than 2*strip (write_sam strips the
 * '<' '>' decoration the same way).  Scoring = calc_sw_score; end cell = first maximum in row-major order; traceback
 * prefers diagonal, then D, then I.  Outputs per (read, slot): score, window interval [ref_beg, ref_end), read interval
 * [q_beg, q_end) (0-based, in the orientation of the window: a reverse-strand window aligns the read as given against the
 * reverse complement), n_ops (> max_ops = truncated: the words are then unspecified), nm (mismatches + gap bases) and ops[max_ops] words
 * (length << 4 | op; 0 = M, 1 = I, 2 = D) in alignment order; soft clips are implied by the read interval. */
int drm_sw_align(drm_ref *r, const uint64_t *ids, int64_t n, int per_read, const char *q_flat, const int64_t *q_off, int strip,
                 int max_ops, int32_t *score, int32_t *ref_beg, int32_t *ref_end, int32_t *q_beg, int32_t *q_end, int32_t *n_ops,
                 int32_t *nm, uint32_t *ops);

/* ------------------------------------------------------------------ input side on the device */
/* extract_FASTA_sequence (src/utils/parse_inputs.cpp:174-220) as a device stream compaction: data = the FASTA file's
 * bytes (host); the first line is dropped, every other byte is upper-cased and kept iff it is A, C, G, T or N.
 * *len = cleaned length; out (host, nullable) receives the bytes when cap >= *len (cap >= size always suffices). */
int drm_fasta_clean(const char *data, int64_t size, int device, char *out, int64_t cap, int64_t *len);
/* The same, followed by drm_ref_open_genome on the result: the `ref_genome` of the dynamic mode
 * (src/main.cpp:178-199) straight from the file bytes.  genome_len (nullable) = number of bases. */
int drm_ref_open_fasta(const char *data, int64_t size, int ref_len, const int *devices, int n_dev, drm_ref **out,
                       int64_t *genome_len);
/* format_fastq (src/utils/parse_inputs.cpp:843-950) on the device: data = the FASTQ file's bytes (host); read r is
 * line 4r+1 (lines end at '\n' only), stored as '<' + line + '>' like the reference's query_seqs.  The result lives
 * on every listed device and feeds drm_sw_rerank_device (drm_reads_device) or drm_search_rerank_reads. */
int drm_reads_open_fastq(const char *data, int64_t size, const int *devices, int n_dev, drm_reads **out);
void drm_reads_close(drm_reads *r);
/* info[0..3] = number of reads, bytes of the flat sequence buffer, longest sequence (with decoration), replicas */
int drm_reads_info(const drm_reads *r, int64_t *info);
/* host copies: flat (info[1] bytes, nullable) and off (info[0] + 1 offsets, nullable) */
int drm_reads_copy_out(const drm_reads *r, char *flat, int64_t *off);
int drm_reads_device(const drm_reads *r, int slot, const char **flat_dev, const int64_t **off_dev);
/* drm_search_rerank with the read sequences taken from a drm_reads handle (opened on the same devices as the index):
 * only the embeddings are uploaded.  n <= number of parsed reads; read i pairs with embedding row i. */
int drm_search_rerank_reads(drm_index *ix, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                            const drm_reads *reads, int64_t stride, int64_t k, int64_t k_clusters, float *D, int64_t *I,
                            int32_t *sw_scores, uint64_t *sw_ids, drm_search_stats *stats, double *timing_ms);

/* ------------------------------------------------------------------ embedder (sequence -> 128-d vector) */
/* Vectorizer::vectorize (src/inference/vectorize.cpp:34-141) on the device: the 3-mer tokeniser
 * (Preprocessor::preprocess src/inference/preprocess.cpp:20-42, truncation to max_len = 123 tokens, pad id 0 as
 * prepareBatch src/inference/vectorize.cpp:345-366) and the network of models/finetuned_sgn33-new-a-Apr6.xml
 * (token embedding -> 2 bidirectional GRU layers, hidden 64 -> [forward | backward] final state of layer 2), in fp32.
 * Sequences are tokenised as given - pass them with the '<' ... '>' decoration where the reference does
 * (src/main.cpp:249-257 embeds its decorated query strings; drm_reads holds reads in that form).
 * Where the reference is undefined (fewer than 2 bytes; a token whose hash lands outside its 96-entry table, which
 * takes a byte other than ACGTacgt<>) the token is the pad id 0; other bytes hash as hashToken's arithmetic says
 * (includes/inference/preprocess.hpp:33-52, char2Val = 7).  No CPU fallback: DRM_ECUDA without a device. */
typedef struct drm_embedder drm_embedder;
/* Reads an OpenVINO IR pair the way ov::Core::read_model would (src/inference/fast_model.cpp:3-40). */
int drm_embedder_open_ir(const char *xml_path, const char *bin_path, int device, drm_embedder **out);
/* Weights by pointer (host, fp16 bit patterns as the IR stores them): emb [vocab][64]; W1, R1 [2][192][64]; B1 [2][256];
 * W2 [2][192][128]; R2 [2][192][64]; B2 [2][256]; gate order z, r, h; B = [Wb_z+Rb_z, Wb_r+Rb_r, Wb_h, Rb_h]. */
int drm_embedder_open_raw(const uint16_t *emb, int vocab, const uint16_t *W1, const uint16_t *R1, const uint16_t *B1,
                          const uint16_t *W2, const uint16_t *R2, const uint16_t *B2, int max_len, int device, drm_embedder **out);
void drm_embedder_close(drm_embedder *e);
/* Host-only check of an IR pair (no device needed): info[0] = vocabulary rows, info[1..7] = element counts of
 * emb, W1, R1, B1, W2, R2, B2; sums (nullable, 7 doubles) = the sum of each tensor's values. */
int drm_embedder_ir_info(const char *xml_path, const char *bin_path, int64_t *info, double *sums);
/* Preprocessor::preprocess on the host (the same code the kernel's prologue runs; format-level helper, no device):
 * tokens = n x max_len vocabulary ids, 0-padded. */
int drm_tokenize(const char *seq_flat, const int64_t *seq_off, int64_t n, int max_len, uint16_t *tokens);
/* n byte strings (host; string i = seq_flat[seq_off[i] .. seq_off[i+1])) -> out (host, n x 128 floats). */
int drm_embed(drm_embedder *e, const char *seq_flat, const int64_t *seq_off, int64_t n, float *out);
/* Reads [first, first + n) of a drm_reads handle that has a replica on the embedder's device -> out (host, n x 128). */
int drm_embed_reads(drm_embedder *e, const drm_reads *reads, int64_t first, int64_t n, float *out);
/* Windows first_id, first_id + id_step, ... (n of them; find_sequence ids, src/utils/post_processor.cpp:47-66) of a
 * reference with a replica on the embedder's device, each embedded as '<' + window + '>' like the strings
 * format_fasta builds (src/utils/parse_inputs.cpp:340-343) -> out (host, n x 128).  A window outside the reference
 * embeds as the empty string. */
int drm_embed_windows(drm_embedder *e, const drm_ref *ref, uint64_t first_id, uint64_t id_step, int64_t n, float *out);
/* Device-pointer forms (pointers on the embedder's device, work queued on `stream`): */
int drm_embed_device(drm_embedder *e, const char *seq_flat_dev, const int64_t *seq_off_dev, int64_t first, int64_t n,
                     float *out_dev, void *stream);
int drm_embed_windows_device(drm_embedder *e, const drm_ref *ref, uint64_t first_id, uint64_t id_step, int64_t n,
                             float *out_dev, void *stream);

/* ------------------------------------------------------------------ L2 rerank stage (embedding distance) */
/* post_process_l2_static (mode 0, src/utils/post_processor.cpp:1023-1162) / post_process_l2_dynamic (mode 1, :552-748)
 * with batch_reranker (src/utils/reranker.cpp:98-196): the candidates of every query are expanded from its search
 * labels, their windows embedded on the device (de-duplicated), ranked by calc_l2_dist (src/utils/metrics.cpp:47-61) to
 * the query's embedding with std::partial_sort's order.  neighbors / distances: n x k_in search output (host); q_emb:
 * n x 128 (host).  Outputs are FLAT like the reference's final vectors: k values per query in the dynamic variant,
 * k_clusters in the static one (what the reference passes to batch_reranker), queries without candidates contribute
 * nothing; capacity n * max(k, k_clusters); *out_count = values written.  On a dense index (stride 1) the dynamic
 * variant passes the first k search results through, like the reference.  The reference's flat slice arithmetic is kept
 * (DESIGN.md); inputs for which it would read past its candidate list are refused with DRM_EINVAL.
 * mode 2 = the static variant over BARE windows: the reference's own callers hand post_process_l2_static the table
 * read_file(ref_seqs_file, ref_len, 1, lookup_mode = true) returns (src/main.cpp:190, src/test_postprocess.cpp:139), whose
 * strings carry no '<' '>' (src/utils/parse_inputs.cpp:345-349), so candidates are embedded undecorated.
 * Window cache: when a call's flat candidate list is at least as long as the reference has windows (DRM_L2_CACHE_SHARE x), and
 * all of the reference's window embeddings fit DRM_L2_CACHE_GB (default 16 GB; 0 turns the cache off), the call embeds EVERY
 * window of `ref` once into a table that the embedder handle keeps until drm_embedder_close (or until a call on another
 * reference or window form replaces it); later calls on the same reference embed nothing.  Same results either way.  One
 * handle serves one call at a time.
 * STATUS: on B200 the device result equals the reference's own post_process_l2_dynamic compiled unmodified (tools/l2_bench.py);
 * 500 k reads/s with the cache warm, 44 k on a first call. */
int drm_l2_rerank(drm_embedder *e, drm_ref *ref, int mode, const int64_t *neighbors, const float *distances, int64_t n, int64_t k_in,
                  const float *q_emb, int64_t stride, int64_t k, int64_t k_clusters, float *out_scores, uint64_t *out_ids,
                  int64_t *out_count);
/* Stage times of this thread's last drm_l2_rerank call: ms5 = {candidate plan on the host, embedder kernels, distance +
 * selection kernels, copies and host bookkeeping, number of query batches}.  Measurement aid (tools/l2_bench.py). */
int drm_l2_last_timing(double *ms5);
/* Host-only pieces of the above, exposed for tests: the candidate plan (info[0..4] = unique windows, flat candidates,
 * results per query, pass-through flag, pass-through count; arrays nullable, sized from a first call) and the
 * partial_sort order (idx: n entries, the first k are the selection).  mode -1 / -2 = the dynamic / static plan with the
 * window ids themselves in `map` (what drm_l2_rerank uses when the window cache applies). */
int drm_l2_plan_host(int mode, const int64_t *neighbors, int64_t n, int64_t k_in, int64_t stride, int64_t k, int64_t k_clusters,
                     uint64_t bound, int64_t *info, uint64_t *unique, uint32_t *map, int64_t *starts);
int drm_l2_select_host(const float *dist, int64_t n, int64_t k, int64_t *idx);

/* ------------------------------------------------------------------ sharded-index merge */
/* D_all/I_all: n_shards x n x k (the all-gathered per-shard results, each row ascending, I already
 * offset to global ids); writes the global top-k per read to D_out/I_out (n x k), ordered by (D, I). */
int drm_merge_topk_device(int device, const float *D_all, const int64_t *I_all, int n_shards, int64_t n, int k,
                          float *D_out, int64_t *I_out, void *stream);

/* ------------------------------------------------------------------ position-sharded index (one HNSW-PQ graph per GPU) */
/* For a reference too large for one GPU (BASELINE.json configs[4]; SURVEY.md §8b `shard_mode = 1`, §8e): the windows are
 * cut by position into W shards, each its own IHNp file with labels local to the shard (id_lo = the shard's first global
 * window id).  One call stands where the reference makes one faiss_search call on one index (src/main.cpp:236-237,278):
 * every read is searched on every shard, the per-shard top-k lists meet on the GPU that owns the read (each chunk of reads
 * is dealt out in W contiguous slices), which merges them by (distance, id) and - in drm_shards_search_rerank - SW-reranks
 * its slice from its replica of the reference windows (post_process_sw_*, src/utils/post_processor.cpp:357-549).
 * A sharded search explores W graphs with ef each, so it is NOT result-identical to one big graph; it equals the CPU
 * oracle run the same sharded way.  The three exchange modes return identical bits:                                    */
#define DRM_XCHG_ALLGATHER 0 /* ncclAllGather of every rank's lists, then merge the slice: W x chunk x k x 12 B per GPU   */
#define DRM_XCHG_ALLTOALL 1  /* grouped ncclSend/ncclRecv: a GPU receives only the lists it merges, chunk x k x 12 B      */
#define DRM_XCHG_FUSED 2     /* the search kernel writes each list into its owner's merge buffer (peer memory over NVLink)
                              * as the read finishes; the only collective left is a 4-byte all-reduce per chunk (barrier). */
typedef struct drm_shards drm_shards;
typedef struct drm_shard_stats {
    double search_ms, exchange_ms, merge_ms, rerank_ms; /* device time of this rank's stages, summed over its steps        */
    int64_t nhops0, nhops_up, ndis_eval;               /* this shard's search counters                                     */
    int64_t bytes_received;                            /* list bytes that arrived from OTHER ranks                         */
    int64_t steps;
} drm_shard_stats;

/* One process drives all W GPUs (one host thread per GPU inside the calls): shard g = index_files[g] on devices[g] (all
 * distinct).  NCCL (ncclCommInitAll) is loaded with dlopen at this point; DRM_EUNSUPPORTED if W > 1 and it is missing.   */
int drm_shards_open(const char *const *index_files, const int64_t *id_lo, const int *devices, int n_shards, drm_shards **out);
/* One process per GPU (torchrun, mpirun): rank 0 obtains the 128-byte ncclUniqueId, the launcher's own channel carries it
 * to the other ranks, every rank opens its shard.                                                                         */
int drm_shards_unique_id(void *id128);
int drm_shards_open_rank(const char *index_file, int64_t id_lo, int device, int rank, int world, const void *id128, drm_shards **out);
void drm_shards_close(drm_shards *h);
/* info[0..8] = world, local ranks, one-process model, peer mapping usable, local vectors, local image bytes, first local
 * rank, its id_lo, NCCL version code                                                                                       */
int drm_shards_info(const drm_shards *h, int64_t *info);
/* HOST buffers.  q: n x d embeddings (every rank reads all of them); D / I: n x k_search global top-k (nullable); with a
 * reference handle that has a replica on every shard's device and non-NULL sw_* the SW rerank runs on the merged labels
 * exactly as drm_sw_rerank would (sw_scores / sw_ids: n x k).  In the one-process model all rows are written; in the
 * rank-per-process model only this rank's rows (slices of every chunk) are.  chunk = reads per exchange step (0 = 262144). */
int drm_shards_search_rerank(drm_shards *h, drm_ref *ref, int mode, const float *q, int64_t n, int d, int k_search, int ef,
                             const char *q_flat, const int64_t *q_off, int64_t stride, int64_t k, int64_t k_clusters,
                             int exchange, int64_t chunk, float *D, int64_t *I, int32_t *sw_scores, uint64_t *sw_ids,
                             drm_shard_stats *stats);
/* DEVICE buffers, one chunk, one local rank (rank-per-process model): q_dev = the chunk's n x d embeddings on this rank's
 * GPU.  Enqueues search -> exchange -> merge on the handle's stream (*step_stream) and returns the merged lists of this
 * rank's slice [slice_lo, slice_lo + slice_n) of the chunk: device pointers owned by the handle, valid until the
 * next-but-one step.  The caller orders its own work after *step_stream.                                                  */
int drm_shards_step_device(drm_shards *h, int local_rank, const float *q_dev, int64_t n, int d, int k, int ef, int exchange,
                           void *reserved, const float **D_merged, const int64_t **I_merged, int64_t *slice_lo,
                           int64_t *slice_n, void **step_stream);
/* Synchronises the rank's stream and reports (optionally resets) its accumulated stage times and counters.  Stage times are
 * accumulated by drm_shards_search_rerank; after drm_shards_step_device time the stages with events on *step_stream.       */
int drm_shards_stats(drm_shards *h, int local_rank, int reset, drm_shard_stats *st);

/* ------------------------------------------------------------------ index build (host, OpenMP) */
/* build_faiss_index (src/hnswpq/index.cpp:86-193): train the product quantizer on `train`
 * (n_train x d; NULL = use x), encode x (n x d), build the HNSW graph with efConstruction = efc and
 * write a FAISS IHNp file.  Not part of the timed search path. */
int drm_build_index(const float *x, int64_t n, int d, const float *train, int64_t n_train, int M_pq, int nbits,
                    int M_hnsw, int efc, int threads, uint32_t seed, const char *out_path);

/* ------------------------------------------------------------------ index build (GPU) */
/* The same job on the device (csrc/build_gpu.cu): PQ k-means + encoding kernels and batched HNSW insertion
 * (one warp per new vertex searches the frozen graph with efConstruction = efc and selects its neighbours with
 * FAISS's shrink_neighbor_list rule; back links are applied per batch under per-vertex locks).  Writes the same
 * FAISS IHNp file.  x/train are HOST pointers (train NULL = use x).  timing_s (nullable): 5 doubles = seconds of
 * upload, PQ training, encoding, graph construction, export.  M_hnsw <= 16, M_pq in {8, 16}, nbits == 8.
 * No CPU fallback: DRM_ECUDA without a device. */
int drm_build_index_gpu(const float *x, int64_t n, int d, const float *train, int64_t n_train, int M_pq, int nbits,
                        int M_hnsw, int efc, int device, uint32_t seed, const char *out_path, double *timing_s);
/* Same with the vectors already resident on `device` (x_dev: n x d fp32 device pointer); the training sample is rows
 * 0, train_step, 2*train_step, ... (the reference's evenly strided sample, src/hnswpq/index.cpp:57-84). */
int drm_build_index_gpu_device(const float *x_dev, int64_t n, int d, int64_t train_step, int M_pq, int nbits, int M_hnsw,
                               int efc, int device, uint32_t seed, const char *out_path, double *timing_s);

#ifdef __cplusplus
}
#endif
#endif /* DRM_B200_H */
