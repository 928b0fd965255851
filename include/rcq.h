/*
 * rcq.h — C ABI of librnaclique_b200.so, the B200-native replacement for RNA-clique's
 * all-vs-all hot path.  Every entry point cites the reference interface it replaces
 * (paths are under the upstream repository, actapia/rna_clique @ a55e40a).
 *
 * Conventions: every function returns 0 on success or a negative rcq_status; the message
 * of the last failure on the calling thread is returned by rcq_last_error().  Nothing
 * throws across the boundary.  Host buffers are caller-owned; handles (rcq_ctx, rcq_sample,
 * rcq_table) are library-owned device objects released with the matching *_free call.
 * One host thread per context.  No torch types appear in any signature.
 */
#ifndef RCQ_H
#define RCQ_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    RCQ_OK = 0,
    RCQ_ERR_CUDA = -1,      /* CUDA runtime failure (message has the CUDA error string) */
    RCQ_ERR_ARG = -2,       /* bad argument */
    RCQ_ERR_CAPACITY = -3,  /* an internal work list overflowed its hard cap */
    RCQ_ERR_NO_IDEAL = -4   /* phase 2: some table restricted to ideal components is empty
                               (reference: NoIdealComponentsError, filtered_distance.py:242-247) */
} rcq_status;

typedef struct rcq_ctx rcq_ctx;
typedef struct rcq_sample rcq_sample;
typedef struct rcq_table rcq_table;

const char* rcq_last_error(void);
int rcq_version(void);

/* One context = one GPU + one stream + scratch arenas.  Replaces nothing in the reference
 * (which has no device); plays the role of the joblib worker pool of find_all_pairs.py:224. */
int rcq_ctx_create(int device, rcq_ctx** out);
void rcq_ctx_destroy(rcq_ctx* ctx);
int rcq_ctx_sync(rcq_ctx* ctx);
/* cudaStream_t of the context, as void* (for timing with events on the launching stream). */
void* rcq_ctx_stream(rcq_ctx* ctx);
/* number of kernels this library launched on ctx since creation (bench.py's gpu_launches) */
int64_t rcq_ctx_launches(rcq_ctx* ctx);
/* device time per stage, measured with CUDA events on the context's stream and accumulated over
 * calls: 0 pack+index, 1 seed join, 2 ungapped, 3 sort/group, 4 gapped DP, 5 selection+table,
 * 6 phase 2, 7 seed probe alone (also inside 1).  reset != 0 clears the accumulators after reading. */
int rcq_ctx_stage_times(rcq_ctx* ctx, double ms[8], int64_t calls[8], int reset);

/* ---- top-gene selection (host only, no device work).  Replaces TopGeneSelector.get_top_genes /
 * get_top_gene_seqs (select_top_genes.py:108-127): per gene the maximum isoform coverage (starting
 * from 0.0), the `top` genes by (coverage, gene id) descending — ties go to the larger gene id, as
 * heapq.nlargest over tuples does — and keep[i] = 1 for every transcript of a selected gene (file
 * order is the caller's).  Returns the number of kept transcripts, or a negative rcq_status. */
int64_t rcq_top_gene_mask(const double* coverage, const int32_t* gene, int64_t n, int64_t top,
                          uint8_t* keep);

/* ---- FASTA ingest (host only).  Replaces the two Bio.SeqIO.parse passes of select_top_genes.py:110-127
 * and, for the default rnaSPAdes pattern (transcripts.py:8), the per-record regex of transcripts.py:69-126.
 * rcq_fasta_scan counts the records and bounds the letters from above (file size minus header lines; line ends
 * inside the sequence lines are still counted) by hopping between '>' characters; rcq_fasta_parse writes the
 * letters (newlines, CR and spaces removed) into ascii_out[], offsets[n_records+1] (offsets[n_records] = the true
 * letter count) and the byte span of every header line;
 * rcq_parse_spades_ids decodes (coverage, gene, isoform) of every record id (first whitespace-separated
 * token of the header) and returns -1, or the index of the first id the pattern does not match. */
int rcq_fasta_scan(const uint8_t* raw, int64_t n, int64_t* n_records, int64_t* n_letters);
int rcq_fasta_parse(const uint8_t* raw, int64_t n, uint8_t* ascii_out, int64_t* offsets,
                    int64_t* hdr_start, int64_t* hdr_end);
/* Lean ingest for a caller that keeps only some records (the top-gene selection of select_top_genes.py:108-127
 * needs the ids of all records but the letters of the selected ones): rcq_fasta_headers = header spans of every
 * record without touching the sequence bytes (n_records from rcq_fasta_scan); rcq_fasta_extract = the letters
 * (LF, CR, spaces removed) of the records keep[0 .. n_keep) concatenated into out[cap], out_off[n_keep + 1] their
 * offsets; returns the number of letters, or -1 with the error set. */
int rcq_fasta_headers(const uint8_t* raw, int64_t n, int64_t n_records, int64_t* hdr_start, int64_t* hdr_end);
int64_t rcq_fasta_extract(const uint8_t* raw, int64_t n, const int64_t* hdr_start, const int64_t* hdr_end,
                          int64_t n_records, const int64_t* keep, int64_t n_keep, uint8_t* out, int64_t cap,
                          int64_t* out_off);
/* FASTA text of selected records ('>' header LF, `width` letters per line, as Bio.SeqIO.write does for
 * od1/<sample>_top.fasta, select_top_genes_all.py:37-45); out == NULL returns the size only. */
int64_t rcq_fasta_format(const uint8_t* ascii, const int64_t* starts, const int64_t* lengths, int64_t n,
                         const uint8_t* hdr, const int64_t* hdr_off, int32_t width, uint8_t* out);
/* The same text written to a file straight from the mapped input (header spans) and the letters rcq_fasta_extract
 * produced, streamed through a small buffer; returns the bytes written, or -1 with the error set. */
int64_t rcq_fasta_write(const char* path, const uint8_t* raw, const int64_t* hdr_start, const int64_t* hdr_end,
                        int64_t n_records, const int64_t* keep, int64_t n_keep, const uint8_t* letters,
                        const int64_t* off, int32_t width);
int64_t rcq_parse_spades_ids(const uint8_t* raw, const int64_t* hdr_start, const int64_t* hdr_end,
                             int64_t n_records, double* coverage, int32_t* gene, int32_t* iso);

/* ---- graph.pkl (host only).  Replaces build_graph (build_graph.py:40-68) + pickle.dump(graph, f,
 * HIGHEST_PROTOCOL) (filtering_step.py:158-159) for the on-disk contract: writes the pickle stream of a
 * plain networkx.Graph whose nodes are (sample_names[s], gene) tuples, one undirected edge per distinct
 * (sgene, qgene) row of the groups (group g = rows [g_off[g], g_off[g+1]) between samples g_s[g] and
 * g_q[g]), nodes in first-appearance order, neighbours in insertion order, plus isolated nodes.
 * counts[0] = nodes, [1] = edges, [2] = bytes written. */
int rcq_graph_pickle(const char* path, int32_t n_samples, const char* const* sample_names,
                     int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_off,
                     const int32_t* sgene, const int32_t* qgene, int64_t n_isolated,
                     const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]);

/* The same writer fed group by group (no concatenated copy of the columns): group g has g_len[g] rows in
 * sgene[g][] / qgene[g][]. */
int rcq_graph_pickle_groups(const char* path, int32_t n_samples, const char* const* sample_names,
                            int32_t n_groups, const int32_t* g_s, const int32_t* g_q, const int64_t* g_len,
                            const int32_t* const* sgene, const int32_t* const* qgene, int64_t n_isolated,
                            const int32_t* iso_sample, const int32_t* iso_gene, int64_t counts[3]);

/* graph.pkl from DEVICE-RESIDENT tables (csrc/graph_device.cu): the same file as rcq_graph_pickle_groups on the
 * tables' (sgene, qgene) columns, with node numbering, duplicate rows, adjacency lists and memo numbers computed on
 * the GPU and only the byte stream produced on the host.  tables[t] joins samples t_s[t] (sgene) and t_q[t]
 * (qgene).  Returns RCQ_ERR_ARG when a table is not ordered by (qgene, sgene) or two tables join the same two
 * samples, RCQ_ERR_CAPACITY for 2^31 rows or gene ids too sparse for dense tables: use the host writer then.
 * The arrays reach the host through ctx's page-locked staging, which grows to 16 bytes per node + 8 bytes per
 * adjacency entry (0.35 GB at BASELINE configs[2]) and is kept until rcq_ctx_destroy. */
int rcq_graph_pickle_tables(rcq_ctx* ctx, const char* path, int32_t n_samples, const char* const* sample_names,
                            int32_t n_tables, const rcq_table* const* tables, const int32_t* t_s, const int32_t* t_q,
                            int64_t counts[3]);

/* ---- matrix.h5 (host only).  Replaces `mat.to_hdf(output_matrix, key="matrix")` (rna_clique.py:175-177; format:
 * docs/formats.md:346-352, read back by export_matrix.py:113 pd.read_hdf): writes an HDF5 file holding ONE pandas
 * fixed-format DataFrame under `key` — group attributes pandas_type="frame", pandas_version, encoding, errors,
 * ndim, nblocks, *_variety; arrays axis0 (column labels), axis1 (row labels), block0_items, block0_values
 * (float64, n_cols x n_rows) — without PyTables or libhdf5: superblock version 0, version-1 object headers,
 * symbol-table groups, contiguous layout.  values is row-major n_rows x n_cols; labels are UTF-8 strings. */
int rcq_matrix_h5_write(const char* path, const char* key, int64_t n_rows, int64_t n_cols,
                        const char* const* row_labels, const char* const* col_labels, const double* values);

/* ---- kernels (1)+(2): pack + seed index.  Replaces makeblastdb per sample
 * (find_all_pairs.py:120-159 make_one_db/make_all_dbs -> BlastDBCache.makedb).
 * ascii[ascii_len]: nucleotide letters (no newlines) holding at least the selected transcripts;
 * transcript t is ascii[starts[t] .. starts[t]+lengths[t]) — transcripts need not be contiguous,
 * so a top-gene selection is passed as (starts, lengths) without gathering.  gene/iso: parsed
 * FASTA ids (transcripts.py:60-126).  ascii_on_device != 0 means `ascii` is a device pointer. */
int rcq_sample_create(rcq_ctx* ctx, const uint8_t* ascii, int64_t ascii_len, int ascii_on_device,
                      const int64_t* starts, const int32_t* lengths, const int32_t* gene,
                      const int32_t* iso, int32_t n_seq, rcq_sample** out);
void rcq_sample_free(rcq_sample* s);
/* info[0]=n_seq, [1]=total bases, [2]=arena bases (padded), [3]=#index entries, [4]=#masked bases */
int rcq_sample_info(const rcq_sample* s, int64_t info[5]);
/* test hooks: copy the packed arenas / index back.  Any pointer may be NULL.
 * fwd/rc: 2-bit words (arena_bases/16 uint32 each); mask: arena_bases/32 uint32;
 * arena_off: int64[n_seq]; index: uint64[#entries] = hash<<32 | pos<<1 | orient */
int rcq_sample_download(const rcq_sample* s, uint32_t* fwd, uint32_t* rc, uint32_t* mask,
                        int64_t* arena_off, uint64_t* index);

/* Low-complexity filter.  blastn soft-masks every query with symmetric DUST (default "-dust 20 64 1") before
 * the search the reference starts at find_homologs.py:124: masked stretches (poly-A tails, microsatellites)
 * take no part in seeding, extensions run through them.  rcq_sample_create computes the same mask on the
 * device (csrc/dust_kernels.cuh) into the sample's seeding mask; one search serves both blastn directions, so
 * a seed needs a 28-base window clean in BOTH samples.  rcq_ctx_set_dust(ctx, 0) switches the filter off for
 * samples created afterwards on ctx (blastn -dust no); returns the previous setting.
 * rcq_sample_seed_mask (test hook): the seeding mask = not-ACGT | DUST | padding, arena_bases/32 uint32, and
 * the number of bases only DUST masks.  Either pointer may be NULL. */
int rcq_ctx_set_dust(rcq_ctx* ctx, int on);
int rcq_sample_seed_mask(const rcq_sample* s, uint32_t* smask, int64_t* n_dust);

/* The same DUST seeding mask without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Runs the kernels rcq_sample_create runs for the mask (csrc/dust_kernels.cuh) with
 * their scratch — one triplet byte per base, one flag bit per base, counters — in a file-scope
 * `static __device__` array of the library (18.1 MB, held by every GPU that loads the library until the process
 * exits) instead of the context's memory pool; the call itself performs no cudaMalloc / cudaFree and needs no
 * rcq_ctx.  Inputs and output are the caller's DEVICE buffers on `device`, laid out as rcq_sample_download /
 * rcq_sample_seed_mask return them: dev_fwd = 2-bit words (arena_bases / 16 uint32), dev_mask = not-ACGT and
 * padding bits (arena_bases / 32 uint32), dev_smask = output (arena_bases / 32 uint32) = dev_mask | DUST.
 * arena_bases: a positive multiple of 64, at most RCQ_DUST_STATIC_MAX_BASES — larger inputs are rejected with
 * RCQ_ERR_CAPACITY and a message naming the limit.  stream: a cudaStream_t (NULL = the default stream); the work
 * is queued on it and the call returns without waiting unless n_dust (host, may be NULL: bases masked by DUST
 * alone) is asked for.  All calls on a GPU share the one scratch array, which keeps the previous call's
 * contents (what has to start at zero is cleared by the call): calls that OVERLAP on the same GPU — two host
 * threads, or two streams without an ordering between them — race; serialise them. */
#define RCQ_DUST_STATIC_MAX_BASES (1ll << 24)
int rcq_dust_mask_static(int device, const uint32_t* dev_fwd, const uint32_t* dev_mask, int64_t arena_bases,
                         uint32_t* dev_smask, void* stream, int64_t* n_dust);

/* ---- kernel (3) + table assembly.  One UNORDERED sample pair: both blastn directions
 * (find_homologs.py:82-133 gene_matches, :124 TabularBlastnSearch), the per-query-gene
 * best-bitscore reduction (:135-164) and HomologFinder.get_match_table (:215-302).
 * s1 plays transcripts1 (ssample side), s2 transcripts2 (qsample side). */
typedef struct {
    int64_t n_candidates;   /* minimizer join key matches examined by the probe */
    int64_t n_seeds;        /* exact runs >= 28 that triggered an ungapped extension */
    int64_t n_ungapped;     /* distinct ungapped HSPs */
    int64_t n_gapped;       /* gapped HSPs produced */
    int64_t dp_cells;       /* DP cells evaluated by the gapped kernel */
    int64_t ungapped_cells; /* base comparisons of the ungapped stage */
    int64_t n_fwd, n_bwd;   /* rows of the two direction tables after best-hit selection */
    int64_t n_rows;         /* rows of the gene-matches table */
    int64_t band_retries;   /* gapped extensions re-run with a wider band */
    int64_t n_verified;     /* key matches that survived the flank filter and were verified on the sequences */
} rcq_search_stats;

/* top_n >= 1: rows per query gene kept in each direction before the reciprocal filter (the reference's
 * top_matches, rna_clique.py:37; nlargest(n, keep="all"), find_homologs.py:135-164).  keep_all = 0 keeps the
 * first best row per query gene only (keep="first").  On error *out is NULL and nothing is left allocated.
 * Limits (RCQ_ERR_CAPACITY, with a message): < 2^20 genes and < 2^21 transcripts per sample, a sample's
 * arena < 2^31 bases, < 2^31 seed candidates and < 2^22 gapped HSPs per pair. */
int rcq_search_pair(rcq_ctx* ctx, const rcq_sample* s1, const rcq_sample* s2, double evalue,
                    int32_t top_n, int32_t keep_all, rcq_table** out, rcq_search_stats* stats);

/* Gene-matches table columns (docs/formats.md:231-252), structure of arrays. */
typedef struct {
    int32_t *length, *mismatch, *gapopen, *qstart, *qend, *sstart, *send;
    int32_t *score;         /* raw alignment score; bitscore/evalue/pident derive from it on the host */
    int32_t *gaps, *nident;
    int32_t *qgene, *qiso, *sgene, *siso;
    int32_t *qlen;          /* length of the BLAST query transcript of that row (for evalue) */
    uint8_t *minus;         /* sstrand == "minus" */
    uint8_t *reverse;       /* row came from the reverse search (find_homologs.py:239,246) */
} rcq_table_cols;
int64_t rcq_table_rows(const rcq_table* t);
int rcq_table_download(const rcq_table* t, rcq_table_cols* host_cols);
void rcq_table_free(rcq_table* t);
/* test hook: the symmetric gapped-HSP set of the last rcq_search_pair on ctx, 12 int32 per HSP:
 * tx1, tx2, minus, q0, q1, s0, s1 (0-based half-open, s on the strand given), score2 (doubled
 * score), mismatch, gaps, gapopen, nident.  Returns the count; copies min(count, cap) rows. */
int64_t rcq_debug_last_hsps(rcq_ctx* ctx, int32_t* out, int64_t cap);
/* the HSP set is kept only after rcq_debug_keep_hsps(ctx, 1) (off by default: it costs a device-wide
 * allocation per search); returns the previous setting.  Besides the tests it serves the isoform
 * orientation search (export_orthologs.py:285-329 blast_pairwise_get_strands), which needs every HSP of a
 * small self-search rather than the per-gene table. */
int rcq_debug_keep_hsps(rcq_ctx* ctx, int on);
/* Karlin-Altschul helpers shared by the host layer (blastn 1/-2 linear-gap parameters):
 * length adjustment and evalue of a raw score for query length m against a db (n bases, nseq). */
int32_t rcq_length_adjustment(int64_t m, int64_t n, int64_t nseq);
/* pident / evalue / bitscore of every row of a table as blastn's tabular output prints them (the reference parses
 * them from that text, find_homologs.py:124-125): host code, one pass over the integer columns of
 * rcq_table_download; forward rows (reverse == 0) use the statistics of the database (bases_fwd, seqs_fwd) =
 * sample 1, backward rows those of sample 2.  ln_k, ln_2: log(0.46), log(2) as the caller computes them. */
int rcq_table_stats(int64_t n, const int32_t* score, const int32_t* qlen, const uint8_t* reverse,
                    const int32_t* nident, const int32_t* length, int64_t bases_fwd, int64_t seqs_fwd,
                    int64_t bases_bwd, int64_t seqs_bwd, double ln_k, double ln_2, double* pident,
                    double* evalue, double* bitscore);
double rcq_evalue(int32_t score, int64_t m, int64_t n, int64_t nseq);
double rcq_bitscore(int32_t score);

/* Measurement helper for bench.py (not on the product path): the INT32 instruction-issue ceiling of the GPU in
 * lane-operations per second (one thread executing one 32-bit integer instruction), from a kernel of
 * independent add / xor / shift chains at full occupancy; best of `reps` launches.  The integer roofline
 * k_gapped is held against. */
int rcq_int32_issue_peak(rcq_ctx* ctx, int32_t reps, double* lane_ops_per_second);

/* ---- kernel (4): gene-match graph, ideal components, per-table reductions.
 * Replaces build_graph.py:40-68, filtered_distance.py:25-39,171-210,234-247 and
 * similarity_computer.py:21-42.  Node id = sample_index*stride + gene.
 * Rows of all tables are concatenated; table k owns rows [row_off[k], row_off[k+1]) and
 * joins samples t_s[k] (ssample) and t_q[k] (qsample).  t_flags (may be NULL = 3 for every table):
 * bit0 = the table's rows are edges of the graph, bit1 = its rows are summed — SampleSimilarity
 * takes the graph and the tables as separate arguments (filtered_distance.py:162-169), so a graph
 * that is not the union of the tables is passed as edge-only pseudo-tables.  sample_count<=0 =>
 * count the samples present in the graph (filtered_distance.py:171-182).
 * Outputs: nident_sum/denom_sum int64[n_tables] (denom = sum(length) - sum(gaps) over
 * restricted rows); valid uint8[n_samples*stride] (may be NULL); counts[0]=#components,
 * [1]=#ideal components, [2]=sample_count used, [3]=#nodes, [4]=#unique edges,
 * [5]=#components with at least sample_count nodes (the --statistics line of plot_component_sizes.py:214-230
 * is counts[2], counts[0], counts[5], counts[1]).
 * Returns RCQ_ERR_NO_IDEAL when some summed table's denom_sum is 0 (sums are still filled in). */
int rcq_phase2_host(rcq_ctx* ctx, int32_t n_samples, int64_t stride, int32_t n_tables,
                    const int64_t* row_off, const int32_t* t_s, const int32_t* t_q,
                    const int32_t* sgene, const int32_t* qgene, const int32_t* nident,
                    const int32_t* length, const int32_t* gaps, const uint8_t* t_flags,
                    int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum, uint8_t* valid,
                    int64_t counts[6]);
/* Same, fed from device-resident tables produced by rcq_search_pair (no host round trip). */
int rcq_phase2_tables(rcq_ctx* ctx, int32_t n_samples, int64_t stride, int32_t n_tables,
                      const rcq_table* const* tables, const int32_t* t_s, const int32_t* t_q,
                      int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum,
                      uint8_t* valid, int64_t counts[6]);

/* ---- multi-GPU: a built sample as one flat device buffer.  With N ranks each rank builds the samples
 * k = rank mod N (kernels 1+2 are not repeated on every GPU) and broadcasts them over NCCL/NVLink; the
 * receivers import the buffer.  meta: [0]=n_seq [1]=total bases [2]=arena bases [3]=#index entries
 * [4]=#masked bases [5]=longest transcript [6]=#genes [7]=bytes of the flat buffer. */
int rcq_sample_meta(const rcq_sample* s, int64_t meta[8]);
int rcq_sample_export(rcq_ctx* ctx, const rcq_sample* s, uint8_t* dev_flat);
int rcq_sample_import(rcq_ctx* ctx, const int64_t meta[8], const uint8_t* dev_flat, rcq_sample** out);

/* ---- multi-GPU exchange of what phase 2 reads (s_node, q_node, nident, length, gaps): 20 bytes per row.
 * rcq_phase2_pack writes the five columns of this rank's tables as one column-major block [5][n_local] into
 * dev_out (device memory, 5 * sum(rows) int32); the host layer all-gathers the blocks (NCCL).
 * rcq_phase2_packed runs phase 2 on the gathered buffer: table t (row_off as in rcq_phase2_host, global pair
 * order) has column c, row x at dev_packed[src_base[t] + c * src_stride[t] + x]. */
int rcq_phase2_pack(rcq_ctx* ctx, int32_t n_tables, const rcq_table* const* tables, int32_t* dev_out);
int rcq_phase2_packed(rcq_ctx* ctx, int32_t n_samples, int64_t stride, int32_t n_tables, const int64_t* row_off,
                      const int32_t* t_s, const int32_t* t_q, const int64_t* src_base, const int64_t* src_stride,
                      const int32_t* dev_packed, int32_t sample_count, int64_t* nident_sum, int64_t* denom_sum,
                      uint8_t* valid, int64_t counts[6]);

/* ---- multi-GPU exchange: serialise a table to / from a flat int32 buffer that the
 * host layer all-gathers with NCCL (torch.distributed) — device pointers in, device pointers out. */
int64_t rcq_table_flat_words(const rcq_table* t);
int rcq_table_flatten(rcq_ctx* ctx, const rcq_table* t, int32_t* dev_out);
int rcq_table_from_flat(rcq_ctx* ctx, const int32_t* dev_in, int64_t n_rows, rcq_table** out);

#ifdef __cplusplus
}
#endif
#endif
