// host_h5.cu — matrix.h5 without PyTables/libhdf5 (host only, no device work).
//
// Replaces `mat.to_hdf(output_matrix, key="matrix")` of the reference (rna_clique.py:175-177; on-disk contract
// docs/formats.md:346-352: a pandas DataFrame in an HDF5 store under the key `matrix`, pandas "fixed" format;
// consumers: export_matrix.py:113 pd.read_hdf).  pandas' fixed format is a PyTables group holding four arrays,
//   axis0          column labels   (fixed-length byte strings, attrs kind = "string", name = pickled None)
//   axis1          row labels
//   block0_items   labels of the columns of block 0 (one float64 block holds the whole matrix)
//   block0_values  float64, one row per column of the block: shape (n_cols, n_rows)
// and the attributes pandas_type = "frame", pandas_version = "0.15.2", encoding = "UTF-8", errors = "strict",
// ndim = 2, nblocks = 1, axis0_variety = axis1_variety = block0_items_variety = "regular".
// (pandas itself stores block0_values transposed with an attribute `transposed = True`; without the
// attribute its reader takes the array as stored, so the array is written in the reader's own orientation.)
//
// The file is a minimal HDF5 1.x file in the original ("earliest") format every HDF5 library reads:
// superblock version 0, version-1 object headers, old-style groups (symbol-table message -> v1 B-tree node ->
// symbol table node, names in a local heap), contiguous dataset layout (layout message version 3), version-1
// dataspace / datatype / attribute messages.  PyTables' own node attributes (CLASS, VERSION, FLAVOR,
// PYTABLES_FORMAT_VERSION) are written so that the file opens as a native PyTables file.
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "common.cuh"

namespace {

const uint64_t UNDEF = 0xFFFFFFFFFFFFFFFFull;

struct Buf {
    std::vector<uint8_t> b;
    size_t size() const { return b.size(); }
    void u8(uint8_t v) { b.push_back(v); }
    void u16(uint16_t v) { for (int i = 0; i < 2; i++) b.push_back((uint8_t)(v >> (8 * i))); }
    void u32(uint32_t v) { for (int i = 0; i < 4; i++) b.push_back((uint8_t)(v >> (8 * i))); }
    void u64(uint64_t v) { for (int i = 0; i < 8; i++) b.push_back((uint8_t)(v >> (8 * i))); }
    void raw(const void* p, size_t n) { const uint8_t* q = (const uint8_t*)p; b.insert(b.end(), q, q + n); }
    void zeros(size_t n) { b.insert(b.end(), n, 0); }
    void pad8() { while (b.size() & 7) b.push_back(0); }
    void put64(size_t at, uint64_t v) { for (int i = 0; i < 8; i++) b[at + i] = (uint8_t)(v >> (8 * i)); }
    void put32(size_t at, uint32_t v) { for (int i = 0; i < 4; i++) b[at + i] = (uint8_t)(v >> (8 * i)); }
    void put16(size_t at, uint16_t v) { for (int i = 0; i < 2; i++) b[at + i] = (uint8_t)(v >> (8 * i)); }
};

// ---- datatype messages (version 1) -------------------------------------------------------------------
void dt_string(Buf& o, uint32_t size, bool utf8) {         // class 3: fixed-length string, null-terminated padding
    o.u8(0x13); o.u8(utf8 ? 0x10 : 0x00); o.u8(0); o.u8(0); o.u32(size);
}
void dt_int64(Buf& o) {                                    // class 0: little-endian signed 64-bit
    o.u8(0x10); o.u8(0x08); o.u8(0); o.u8(0); o.u32(8); o.u16(0); o.u16(64);
}
void dt_float64(Buf& o) {                                  // class 1: IEEE 754 binary64, little-endian
    o.u8(0x11); o.u8(0x20); o.u8(63); o.u8(0); o.u32(8);
    o.u16(0); o.u16(64); o.u8(52); o.u8(11); o.u8(0); o.u8(52); o.u32(1023);
}
void ds_simple(Buf& o, int rank, const uint64_t* dims) {   // dataspace message version 1 (rank 0 = scalar)
    o.u8(1); o.u8((uint8_t)rank); o.u8(0); o.zeros(5);
    for (int i = 0; i < rank; i++) o.u64(dims[i]);
}

// ---- one header message: type, size, flags, reserved(3), body padded to 8 --------------------------------
struct Header {
    Buf body;
    int n_msgs = 0;
    void msg(uint16_t type, const Buf& data) {
        Buf d = data;
        d.pad8();
        body.u16(type); body.u16((uint16_t)d.size()); body.u8(0); body.zeros(3);
        body.raw(d.b.data(), d.size());
        n_msgs++;
    }
    // attribute message, version 1: name / datatype / dataspace each padded to 8, then the value
    void attr(const char* name, const Buf& dtype, const Buf& dspace, const void* value, size_t nbytes) {
        Buf a;
        const size_t nl = strlen(name) + 1;
        a.u8(1); a.u8(0); a.u16((uint16_t)nl); a.u16((uint16_t)dtype.size()); a.u16((uint16_t)dspace.size());
        a.raw(name, nl); a.pad8();
        a.raw(dtype.b.data(), dtype.size()); a.pad8();
        a.raw(dspace.b.data(), dspace.size()); a.pad8();
        a.raw(value, nbytes);
        msg(0x000C, a);
    }
    void attr_str(const char* name, const char* value, bool utf8 = true) {   // scalar fixed-length string, as PyTables writes str
        Buf t, s;
        const size_t n = strlen(value);
        dt_string(t, (uint32_t)(n ? n : 1), utf8);
        ds_simple(s, 0, nullptr);
        const char zero = 0;
        attr(name, t, s, n ? value : &zero, n ? n : 1);
    }
    void attr_i64(const char* name, int64_t v) {
        Buf t, s;
        dt_int64(t);
        ds_simple(s, 0, nullptr);
        attr(name, t, s, &v, 8);
    }
    // version-1 object header: 16-byte prefix + messages
    void emit(Buf& file) const {
        file.u8(1); file.u8(0); file.u16((uint16_t)n_msgs); file.u32(1); file.u32((uint32_t)body.size()); file.u32(0);
        file.raw(body.b.data(), body.size());
    }
    size_t bytes() const { return 16 + body.size(); }
};

// ---- old-style group: object header (+ attributes) followed by its B-tree node, local heap and symbol table node
struct Child { std::string name; uint64_t header_addr; bool is_group; uint64_t btree, heap; };

// writes the group at the current end of `f`; children must be sorted by name.  Returns header address and
// fills btree/heap addresses.
uint64_t emit_group(Buf& f, Header& h, const std::vector<Child>& kids, uint64_t* btree_out, uint64_t* heap_out) {
    // local heap data: "" at offset 0, then the names
    Buf names;
    names.zeros(8);
    std::vector<uint64_t> name_off;
    for (const Child& c : kids) {
        name_off.push_back(names.size());
        names.raw(c.name.c_str(), c.name.size() + 1);
        names.pad8();
    }
    Buf stab;            // symbol table message body: B-tree address, heap address (patched below)
    stab.u64(0); stab.u64(0);
    h.msg(0x0011, stab);
    const size_t stab_at = h.body.size() - 16;
    const uint64_t header_addr = f.size();
    const uint64_t btree = header_addr + h.bytes();
    const uint64_t heap = btree + 544;
    const uint64_t heap_data = heap + 32;
    const uint64_t snod = heap_data + names.size();
    h.body.put64(stab_at, btree);
    h.body.put64(stab_at + 8, heap);
    h.emit(f);
    // B-tree v1 node, group type, leaf level, one child (K = 16: 33 keys + 32 children)
    f.raw("TREE", 4); f.u8(0); f.u8(0); f.u16(1); f.u64(UNDEF); f.u64(UNDEF);
    f.u64(0); f.u64(snod); f.u64(name_off.empty() ? 0 : name_off.back());
    f.zeros(544 - 24 - 24);
    // local heap
    f.raw("HEAP", 4); f.u8(0); f.zeros(3); f.u64(names.size()); f.u64(1 /* no free block */); f.u64(heap_data);
    f.raw(names.b.data(), names.size());
    // symbol table node (leaf K = 4: room for 8 entries)
    f.raw("SNOD", 4); f.u8(1); f.u8(0); f.u16((uint16_t)kids.size());
    for (size_t i = 0; i < 8; i++) {
        if (i < kids.size()) {
            const Child& c = kids[i];
            f.u64(name_off[i]); f.u64(c.header_addr); f.u32(c.is_group ? 1 : 0); f.u32(0);
            if (c.is_group) { f.u64(c.btree); f.u64(c.heap); } else f.zeros(16);
        } else f.zeros(40);
    }
    *btree_out = btree; *heap_out = heap;
    return header_addr;
}

// ---- one contiguous dataset: header (dataspace, datatype, fill value, layout, attributes) + data
uint64_t emit_dataset(Buf& f, int rank, const uint64_t* dims, const Buf& dtype, uint32_t elem, Header& attrs,
                      const void* data) {
    Header h;
    Buf sp, fill, lay;
    ds_simple(sp, rank, dims);
    h.msg(0x0001, sp);
    h.msg(0x0003, dtype);
    fill.u8(2); fill.u8(2); fill.u8(2); fill.u8(1); fill.u32(0);   // version 2, allocate late, write if set, default fill value (size 0): what the HDF5 library writes itself
    h.msg(0x0005, fill);
    uint64_t n = elem;
    for (int i = 0; i < rank; i++) n *= dims[i];
    lay.u8(3); lay.u8(1); lay.u64(0); lay.u64(n);            // version 3, contiguous, address patched below
    h.msg(0x0008, lay);
    const size_t lay_at = h.body.size() - 24 + 2;
    h.body.raw(attrs.body.b.data(), attrs.body.size());
    h.n_msgs += attrs.n_msgs;
    const uint64_t addr = f.size();
    h.body.put64(lay_at, addr + h.bytes());
    h.emit(f);
    f.raw(data, n);
    f.pad8();
    return addr;
}

uint64_t emit_label_array(Buf& f, int64_t n, const char* const* labels) {
    size_t w = 1;
    for (int64_t i = 0; i < n; i++) w = std::max(w, strlen(labels[i]));
    std::vector<char> data((size_t)n * w, 0);
    for (int64_t i = 0; i < n; i++) memcpy(&data[(size_t)i * w], labels[i], strlen(labels[i]));
    Buf t;
    dt_string(t, (uint32_t)w, false);
    Header a;
    a.attr_str("CLASS", "ARRAY"); a.attr_str("VERSION", "2.4"); a.attr_str("FLAVOR", "numpy");
    a.attr_str("kind", "string"); a.attr_str("name", "N.", false);      // name = pickle.dumps(None, 0)
    const uint64_t dims[1] = {(uint64_t)n};
    return emit_dataset(f, 1, dims, t, (uint32_t)w, a, data.data());
}

}  // namespace

extern "C" int rcq_matrix_h5_write(const char* path, const char* key, int64_t n_rows, int64_t n_cols,
                                   const char* const* row_labels, const char* const* col_labels,
                                   const double* values) {
    if (!path || !key || !*key || n_rows < 1 || n_cols < 1 || !row_labels || !col_labels || !values)
        return rcq_fail(RCQ_ERR_ARG, "rcq_matrix_h5_write: bad argument (an empty frame has no fixed-format layout here)");
    if (strchr(key, '/')) return rcq_fail(RCQ_ERR_ARG, "rcq_matrix_h5_write: the key must be a single path component");
    Buf f;
    f.zeros(96);                                             // superblock, filled in last
    // datasets first (children of the key group), in name order
    std::vector<Child> kids;
    kids.push_back({"axis0", emit_label_array(f, n_cols, col_labels), false, 0, 0});
    kids.push_back({"axis1", emit_label_array(f, n_rows, row_labels), false, 0, 0});
    kids.push_back({"block0_items", emit_label_array(f, n_cols, col_labels), false, 0, 0});
    {
        // pandas reads block values as (items of the block, rows): column-major copy of the row-major matrix
        std::vector<double> t((size_t)n_rows * n_cols);
        for (int64_t r = 0; r < n_rows; r++)
            for (int64_t c = 0; c < n_cols; c++) t[(size_t)c * n_rows + r] = values[(size_t)r * n_cols + c];
        Buf dt;
        dt_float64(dt);
        Header a;
        a.attr_str("CLASS", "ARRAY"); a.attr_str("VERSION", "2.4"); a.attr_str("FLAVOR", "numpy");
        const uint64_t dims[2] = {(uint64_t)n_cols, (uint64_t)n_rows};
        kids.push_back({"block0_values", emit_dataset(f, 2, dims, dt, 8, a, t.data()), false, 0, 0});
    }
    Header g;
    g.attr_str("CLASS", "GROUP"); g.attr_str("VERSION", "1.0");
    g.attr_str("pandas_type", "frame"); g.attr_str("pandas_version", "0.15.2");
    g.attr_str("encoding", "UTF-8"); g.attr_str("errors", "strict");
    g.attr_i64("ndim", 2); g.attr_i64("nblocks", 1);
    g.attr_str("axis0_variety", "regular"); g.attr_str("axis1_variety", "regular");
    g.attr_str("block0_items_variety", "regular");
    uint64_t gb = 0, gh = 0;
    const uint64_t gaddr = emit_group(f, g, kids, &gb, &gh);
    Header root;
    root.attr_str("CLASS", "GROUP"); root.attr_str("VERSION", "1.0"); root.attr_str("PYTABLES_FORMAT_VERSION", "2.1");
    uint64_t rb = 0, rh = 0;
    std::vector<Child> top;
    top.push_back({key, gaddr, true, gb, gh});
    const uint64_t raddr = emit_group(f, root, top, &rb, &rh);
    // superblock version 0
    Buf sb;
    const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    sb.raw(sig, 8);
    sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(8); sb.u8(8); sb.u8(0);
    sb.u16(4); sb.u16(16); sb.u32(0);
    sb.u64(0); sb.u64(UNDEF); sb.u64(f.size()); sb.u64(UNDEF);
    sb.u64(0); sb.u64(raddr); sb.u32(1); sb.u32(0); sb.u64(rb); sb.u64(rh);
    memcpy(f.b.data(), sb.b.data(), 96);
    FILE* fp = fopen(path, "wb");
    if (!fp) return rcq_fail(RCQ_ERR_ARG, "rcq_matrix_h5_write: cannot open %s", path);
    const size_t wrote = fwrite(f.b.data(), 1, f.size(), fp);
    if (fclose(fp) != 0 || wrote != f.size()) return rcq_fail(RCQ_ERR_ARG, "rcq_matrix_h5_write: short write to %s", path);
    return 0;
}
