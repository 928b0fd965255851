// host_fasta.cu — host-side FASTA ingest and rnaSPAdes id decoding (no device work).
// Replaces the Bio.SeqIO.parse passes of select_top_genes.py:110-127 and the per-record regex of
// transcripts.py:8,69-126 for the DEFAULT id pattern; any other pattern stays in Python.
#include <stdlib.h>
#include <string.h>
#include "common.cuh"

static inline bool is_skip(uint8_t c) { return c == '\n' || c == '\r' || c == ' '; }

// Pass 1: number of records (lines starting with '>') and an upper bound of the sequence letters.  '>' is rare
// outside header lines, so the records are counted by hopping from '>' to '>' with memchr; the letter count
// returned is the file size minus the header lines - every LF, CR or space inside the sequence lines makes the
// true count (offsets[n_records] after rcq_fasta_parse) smaller.
extern "C" int rcq_fasta_scan(const uint8_t* raw, int64_t n, int64_t* n_records, int64_t* n_letters) {
    if (n < 0 || (n > 0 && !raw) || !n_records || !n_letters) return rcq_fail(RCQ_ERR_ARG, "rcq_fasta_scan: bad argument");
    int64_t rec = 0, hdr_bytes = 0, p = 0;
    while (p < n) {
        const uint8_t* gt = (const uint8_t*)memchr(raw + p, '>', (size_t)(n - p));
        if (!gt) break;
        const int64_t at = (int64_t)(gt - raw);
        if (at == 0 || raw[at - 1] == '\n') {
            const uint8_t* nl = (const uint8_t*)memchr(gt, '\n', (size_t)(n - at));
            const int64_t e = nl ? (int64_t)(nl - raw) : n;
            rec++;
            hdr_bytes += e - at + 1;
            p = e + 1;
        } else p = at + 1;
    }
    *n_records = rec; *n_letters = n - hdr_bytes > 0 ? n - hdr_bytes : 0;
    return 0;
}

// Pass 2: letters -> ascii_out (n_letters bytes), offsets[r] = letters before record r's header
// (offsets[n_records] = n_letters), hdr_start/hdr_end[r] = byte span of the header line ('>' .. LF).
extern "C" int rcq_fasta_parse(const uint8_t* raw, int64_t n, uint8_t* ascii_out, int64_t* offsets,
                               int64_t* hdr_start, int64_t* hdr_end) {
    if (n < 0 || (n > 0 && (!raw || !offsets))) return rcq_fail(RCQ_ERR_ARG, "rcq_fasta_parse: bad argument");
    int64_t rec = 0, let = 0, p = 0;
    while (p < n) {
        const uint8_t* nl = (const uint8_t*)memchr(raw + p, '\n', (size_t)(n - p));
        const int64_t e = nl ? (int64_t)(nl - raw) : n;
        if (raw[p] == '>') {
            offsets[rec] = let; hdr_start[rec] = p; hdr_end[rec] = e; rec++;
        } else {
            const uint8_t* a = raw + p;
            const int64_t len = e - p;
            if (!memchr(a, '\r', (size_t)len) && !memchr(a, ' ', (size_t)len)) { memcpy(ascii_out + let, a, (size_t)len); let += len; }
            else for (int64_t i = 0; i < len; i++) if (!is_skip(a[i])) ascii_out[let++] = a[i];
        }
        p = e + 1;
    }
    offsets[rec] = let;
    return 0;
}

// Lean ingest for callers that keep only some records (top-gene selection): the header spans of every record
// without touching the sequence bytes (records are found by hopping from '>' to '>' like rcq_fasta_scan; n_records
// is that call's count) ...
extern "C" int rcq_fasta_headers(const uint8_t* raw, int64_t n, int64_t n_records, int64_t* hdr_start, int64_t* hdr_end) {
    if (n < 0 || n_records < 0 || (n > 0 && !raw) || (n_records && (!hdr_start || !hdr_end)))
        return rcq_fail(RCQ_ERR_ARG, "rcq_fasta_headers: bad argument");
    int64_t rec = 0, p = 0;
    while (p < n) {
        const uint8_t* gt = (const uint8_t*)memchr(raw + p, '>', (size_t)(n - p));
        if (!gt) break;
        const int64_t at = (int64_t)(gt - raw);
        if (at == 0 || raw[at - 1] == '\n') {
            const uint8_t* nl = (const uint8_t*)memchr(gt, '\n', (size_t)(n - at));
            const int64_t e = nl ? (int64_t)(nl - raw) : n;
            if (rec >= n_records) return rcq_fail(RCQ_ERR_ARG, "rcq_fasta_headers: more records than announced");
            hdr_start[rec] = at; hdr_end[rec] = e; rec++;
            p = e + 1;
        } else p = at + 1;
    }
    if (rec != n_records) return rcq_fail(RCQ_ERR_ARG, "rcq_fasta_headers: fewer records than announced");
    return 0;
}
// ... and the letters (LF, CR and spaces removed) of the records keep[0 .. n_keep) only, concatenated into out
// (capacity cap bytes; the sequence bytes of a record are a safe bound of its letters), out_off[k] = letters
// before keep[k], out_off[n_keep] = total.  Returns the total, or -1 (error set) when cap is too small.
extern "C" int64_t rcq_fasta_extract(const uint8_t* raw, int64_t n, const int64_t* hdr_start, const int64_t* hdr_end,
                                     int64_t n_records, const int64_t* keep, int64_t n_keep, uint8_t* out, int64_t cap,
                                     int64_t* out_off) {
    if (n < 0 || n_records < 0 || n_keep < 0 || (n > 0 && !raw) || (n_records && (!hdr_start || !hdr_end)) ||
        (n_keep && (!keep || !out)) || !out_off) {
        rcq_fail(RCQ_ERR_ARG, "rcq_fasta_extract: bad argument");
        return -1;
    }
    int64_t let = 0;
    for (int64_t k = 0; k < n_keep; k++) {
        const int64_t r = keep[k];
        if (r < 0 || r >= n_records) { rcq_fail(RCQ_ERR_ARG, "rcq_fasta_extract: record index out of range"); return -1; }
        out_off[k] = let;
        int64_t p = hdr_end[r] + 1;
        const int64_t end = r + 1 < n_records ? hdr_start[r + 1] : n;
        if (end - p > cap - let) { rcq_fail(RCQ_ERR_CAPACITY, "rcq_fasta_extract: output buffer too small"); return -1; }
        // CR and spaces are looked for once per record (two scans of a few kilobytes), not once per line
        const bool plain = !memchr(raw + p, '\r', (size_t)(end - p)) && !memchr(raw + p, ' ', (size_t)(end - p));
        while (p < end) {
            const uint8_t* nl = (const uint8_t*)memchr(raw + p, '\n', (size_t)(end - p));
            const int64_t e = nl ? (int64_t)(nl - raw) : end;
            const uint8_t* a = raw + p;
            const int64_t len = e - p;
            if (plain) { memcpy(out + let, a, (size_t)len); let += len; }
            else for (int64_t i = 0; i < len; i++) if (!is_skip(a[i])) out[let++] = a[i];
            p = e + 1;
        }
    }
    out_off[n_keep] = let;
    return let;
}

// id = first whitespace-separated token of the header text raw[hdr_start+1 .. hdr_end)
static inline void id_span(const uint8_t* raw, int64_t a, int64_t b, int64_t* s, int64_t* e) {
    auto ws = [](uint8_t c) { return c == ' ' || (c >= 9 && c <= 13) || (c >= 28 && c <= 31); };
    while (a < b && ws(raw[a])) a++;
    int64_t q = a;
    while (q < b && !ws(raw[q])) q++;
    *s = a; *e = q;
}
static inline int64_t digits(const uint8_t* p, const uint8_t* end) {
    int64_t k = 0;
    while (p + k < end && p[k] >= '0' && p[k] <= '9') k++;
    return k;
}
// The default pattern ^.*cov_([0-9]+(?:\.[0-9]+))_g([0-9]+)_i([0-9]+) with re.search semantics: the
// greedy .* makes the LAST position where the rest matches win.  Returns -1 when every id matched,
// else the index of the first id that did not (Python raises TranscriptIDParseError for it).
extern "C" int64_t rcq_parse_spades_ids(const uint8_t* raw, const int64_t* hdr_start, const int64_t* hdr_end,
                                        int64_t n_records, double* coverage, int32_t* gene, int32_t* iso) {
    for (int64_t r = 0; r < n_records; r++) {
        int64_t s, e;
        id_span(raw, hdr_start[r] + 1, hdr_end[r], &s, &e);
        const uint8_t* end = raw + e;
        bool ok = false;
        for (int64_t p = e - 4; p >= s && !ok; p--) {
            if (memcmp(raw + p, "cov_", 4) != 0) continue;
            const uint8_t* q = raw + p + 4;
            const int64_t d1 = digits(q, end);
            if (!d1 || q + d1 >= end || q[d1] != '.') continue;
            const int64_t d2 = digits(q + d1 + 1, end);
            if (!d2) continue;
            const uint8_t* g = q + d1 + 1 + d2;
            // [0-9]+ may give digits back to let "_g" match: it cannot (a digit is never '_'), so one try
            if (g + 2 > end || g[0] != '_' || g[1] != 'g') continue;
            const int64_t d3 = digits(g + 2, end);
            if (!d3) continue;
            const uint8_t* i = g + 2 + d3;
            if (i + 2 > end || i[0] != '_' || i[1] != 'i') continue;
            const int64_t d4 = digits(i + 2, end);
            if (!d4 || d3 > 9 || d4 > 9 || d1 + d2 > 300) continue;
            if (d1 + d2 <= 15) {
                // Clinger's fast path: at most 15 significant digits and a power of ten up to 10^15 are exact doubles,
                // and one IEEE division rounds their quotient correctly - the value strtod (and Python's float) gives
                static const double P10[16] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15};
                int64_t m = 0;
                for (int64_t k = 0; k < d1; k++) m = m * 10 + (q[k] - '0');
                for (int64_t k = 0; k < d2; k++) m = m * 10 + (q[d1 + 1 + k] - '0');
                coverage[r] = (double)m / P10[d2];
            } else {
                char buf[320];
                memcpy(buf, q, (size_t)(d1 + 1 + d2)); buf[d1 + 1 + d2] = 0;
                coverage[r] = strtod(buf, nullptr);
            }
            int64_t gv = 0, iv = 0;
            for (int64_t k = 0; k < d3; k++) gv = gv * 10 + (g[2 + k] - '0');
            for (int64_t k = 0; k < d4; k++) iv = iv * 10 + (i[2 + k] - '0');
            gene[r] = (int32_t)gv; iso[r] = (int32_t)iv;
            ok = true;
        }
        if (!ok) return r;
    }
    return -1;
}

// FASTA text of the selected records, Biopython style: '>' header LF, then `width` letters per line
// (select_top_genes_all.py:37-45 writes od1/<sample>_top.fasta with SeqIO.write).  hdr = all header
// texts concatenated, record r's header = hdr[hdr_off[r] .. hdr_off[r+1]).  With out == NULL only the
// size is returned; otherwise the bytes are written and the size returned.
extern "C" int64_t rcq_fasta_format(const uint8_t* ascii, const int64_t* starts, const int64_t* lengths, int64_t n,
                                    const uint8_t* hdr, const int64_t* hdr_off, int32_t width, uint8_t* out) {
    if (n < 0 || width < 1 || (n > 0 && (!starts || !lengths || !hdr_off))) return rcq_fail(RCQ_ERR_ARG, "rcq_fasta_format: bad argument");
    int64_t at = 0;
    for (int64_t r = 0; r < n; r++) {
        const int64_t hl = hdr_off[r + 1] - hdr_off[r], L = lengths[r];
        if (out) { out[at] = '>'; memcpy(out + at + 1, hdr + hdr_off[r], (size_t)hl); out[at + 1 + hl] = '\n'; }
        at += hl + 2;
        for (int64_t p = 0; p < L; p += width) {
            const int64_t k = L - p < width ? L - p : width;
            if (out) { memcpy(out + at, ascii + starts[r] + p, (size_t)k); out[at + k] = '\n'; }
            at += k + 1;
        }
    }
    return at;
}

// od1/<sample>_top.fasta written straight from the mapped input and the extracted letters (the lean path): record
// keep[k] gets '>' + its header text (raw[hdr_start + 1 .. hdr_end), a trailing CR dropped) + LF and the letters
// letters[off[k] .. off[k + 1]) in lines of `width` - the bytes rcq_fasta_format produces - streamed through a small
// buffer instead of being assembled in memory first.  Returns the bytes written, or -1 with the error set.
extern "C" int64_t rcq_fasta_write(const char* path, const uint8_t* raw, const int64_t* hdr_start, const int64_t* hdr_end,
                                   int64_t n_records, const int64_t* keep, int64_t n_keep, const uint8_t* letters,
                                   const int64_t* off, int32_t width) {
    if (!path || n_keep < 0 || width < 1 || (n_keep && (!raw || !hdr_start || !hdr_end || !keep || !letters || !off))) {
        rcq_fail(RCQ_ERR_ARG, "rcq_fasta_write: bad argument");
        return -1;
    }
    FILE* f = fopen(path, "wb");
    if (!f) { rcq_fail(RCQ_ERR_ARG, "rcq_fasta_write: cannot open the output file"); return -1; }
    const size_t CAP = 4u << 20;
    std::vector<uint8_t> buf(CAP + 65536);
    size_t at = 0;
    int64_t total = 0;
    bool ok = true;
    auto flush = [&]() { if (at) { ok = ok && fwrite(buf.data(), 1, at, f) == at; total += (int64_t)at; at = 0; } };
    for (int64_t k = 0; k < n_keep && ok; k++) {
        const int64_t r = keep[k];
        if (r < 0 || r >= n_records) { fclose(f); rcq_fail(RCQ_ERR_ARG, "rcq_fasta_write: record index out of range"); return -1; }
        int64_t a = hdr_start[r] + 1, b = hdr_end[r];
        if (b > a && raw[b - 1] == '\r') b--;
        if (b < a) b = a;
        const size_t hl = (size_t)(b - a);
        if (hl + 2 > 65536) {                      // a header longer than the slack: written on its own
            flush();
            ok = ok && fputc('>', f) != EOF && fwrite(raw + a, 1, hl, f) == hl && fputc('\n', f) != EOF;
            total += (int64_t)hl + 2;
        } else {
            if (at + hl + 2 > CAP + 65536) flush();
            buf[at++] = '>'; memcpy(buf.data() + at, raw + a, hl); at += hl; buf[at++] = '\n';
        }
        const int64_t L = off[k + 1] - off[k];
        const uint8_t* seq = letters + off[k];
        for (int64_t p = 0; p < L; p += width) {
            const size_t n = (size_t)(L - p < width ? L - p : width);
            if (n + 1 > 65536) {                   // absurd line width: no buffering
                flush();
                ok = ok && fwrite(seq + p, 1, n, f) == n && fputc('\n', f) != EOF;
                total += (int64_t)n + 1;
                continue;
            }
            if (at + n + 1 > CAP) flush();
            memcpy(buf.data() + at, seq + p, n); at += n; buf[at++] = '\n';
        }
    }
    flush();
    const bool closed = fclose(f) == 0;
    if (!ok || !closed) { rcq_fail(RCQ_ERR_ARG, "rcq_fasta_write: short write"); return -1; }
    return total;
}
