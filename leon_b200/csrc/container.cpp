// `.leon` container writer (host only, no CUDA): lays the outputs of leon_compress() out as the HDF5 file that the
// reference's `leon -d` opens.  The reference writes the container through the HDF5 library
// (tools/storage/impl/StorageHDF5.hpp:208-215 H5Fcreate, :271-279 groups, CollectionHDF5Patch.hpp:164-200 datasets,
// :99-114 string attributes; C/Leon.cpp:485-516 groups + infobyte, :581-651 qual_<id>, :653-710 dna_<id>/header_<id>,
// :965-1081 bloom, anchors, blocksizes, metadata).  That library is not part of this product, so this file emits the
// HDF5 1.8 on-disk format directly, in its oldest and most widely readable form:
//   superblock v0 -> root group; groups = object header v1 + symbol-table message -> B-tree v1 (TREE) of symbol
//   nodes (SNOD) + local heap (HEAP) of link names; datasets = object header v1 with dataspace (rank 1), datatype
//   (H5T_STD_U8LE), CONTIGUOUS layout and v1 attribute messages; attributes are variable-length strings whose bytes
//   live in global heap collections (GCOL), exactly the datatype the reference reader asks for (H5T_C_S1, H5T_VARIABLE).
// The reader side only uses H5Lexists/H5Gopen2/H5Dopen2/H5Dget_space/H5Dread(hyperslab)/H5Aopen/H5Aread, all of which
// are layout-agnostic, so contiguous storage replaces the reference's chunked one (no chunk B-trees, no filters).
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/leon_b200.h"

extern "C" void leon_internal_set_error(const char *msg);   // leon_b200.cu: message behind leon_last_error(NULL)

namespace {

constexpr uint64_t H5_UNDEF = ~0ull;
constexpr unsigned GROUP_LEAF_K = 4;        // library defaults (H5F_CRT_SYM_LEAF_DEF / btree K of group nodes)
constexpr unsigned GROUP_INTERNAL_K = 16;
constexpr size_t GCOL_MIN = 4096;           // H5HG_MINSIZE

inline uint64_t pad8(uint64_t n) { return (n + 7) & ~7ull; }

struct Bytes {
    std::vector<uint8_t> b;
    void u8(unsigned v) { b.push_back((uint8_t)v); }
    void u16(unsigned v) { u8(v & 255); u8((v >> 8) & 255); }
    void u32(uint32_t v) { u16(v & 0xffff); u16(v >> 16); }
    void u64(uint64_t v) { u32((uint32_t)v); u32((uint32_t)(v >> 32)); }
    void raw(const void *p, size_t n) { const uint8_t *q = (const uint8_t *)p; if (n) b.insert(b.end(), q, q + n); }
    void zeros(size_t n) { b.insert(b.end(), n, 0); }
    void align8() { zeros((size_t)(pad8(b.size()) - b.size())); }
    size_t size() const { return b.size(); }
};

// One extent of the output file: either bytes owned by the writer or a view of the caller's memory.
struct Piece {
    uint64_t off = 0;
    std::vector<uint8_t> own;
    const uint8_t *ext = nullptr;
    uint64_t len = 0;
};

struct Attr { std::string name, value; };

struct Dataset {
    std::string name;
    const uint8_t *data = nullptr;
    uint64_t len = 0;
    std::vector<Attr> attrs;
};

struct Group {
    std::string name;
    std::vector<std::unique_ptr<Group>> groups;
    std::vector<Dataset> dsets;
    Group *group(const std::string &n) {
        groups.emplace_back(new Group());
        groups.back()->name = n;
        return groups.back().get();
    }
    Dataset &dataset(const std::string &n, const void *p, uint64_t len) {
        dsets.emplace_back();
        dsets.back().name = n;
        dsets.back().data = (const uint8_t *)p;
        dsets.back().len = len;
        return dsets.back();
    }
};

struct Link {            // one symbol-table entry
    std::string name;
    uint64_t ohdr = 0;
    bool is_group = false;
    uint64_t btree = 0, heap = 0;
    uint64_t name_off = 0;
};

class H5Writer {
  public:
    uint64_t eof = 0;
    std::vector<Piece> pieces;

    uint64_t alloc(uint64_t n) {
        uint64_t o = pad8(eof);
        eof = o + n;
        return o;
    }
    void put(uint64_t off, Bytes &&v) {
        Piece p;
        p.off = off;
        p.len = v.b.size();
        p.own = std::move(v.b);
        pieces.push_back(std::move(p));
    }
    void put_ext(uint64_t off, const uint8_t *q, uint64_t n) {
        Piece p;
        p.off = off;
        p.ext = q;
        p.len = n;
        pieces.push_back(std::move(p));
    }

    // ---- global heap: variable-length string payloads -------------------------------------------------------
    struct GCol { uint64_t addr = 0; uint64_t size = 0; Bytes body; unsigned next_idx = 1; };
    GCol cur;
    bool cur_open = false;

    void gcol_close() {
        if (!cur_open) return;
        uint64_t used = cur.body.size();
        uint64_t rest = cur.size - used;
        if (rest >= 16) {                       // object 0 = the free space, its size includes its own header
            cur.body.u16(0); cur.body.u16(0); cur.body.u32(0); cur.body.u64(rest);
            cur.body.zeros((size_t)(rest - 16));
        } else {
            cur.body.zeros((size_t)rest);
        }
        put(cur.addr, std::move(cur.body));
        cur = GCol();
        cur_open = false;
    }
    void gcol_open(uint64_t need) {
        gcol_close();
        cur.size = std::max<uint64_t>(GCOL_MIN, pad8(16 + need));
        cur.addr = alloc(cur.size);
        cur.body.raw("GCOL", 4);
        cur.body.u8(1); cur.body.zeros(3);
        cur.body.u64(cur.size);
        cur.next_idx = 1;
        cur_open = true;
    }
    // Stores the string; returns the 16-byte on-disk vlen element {length, collection address, object index}.
    void vlen_string(const std::string &s, Bytes &elem) {
        uint64_t need = 16 + pad8(s.size());
        if (!cur_open || cur.body.size() + need > cur.size || cur.next_idx > 60000) gcol_open(need);
        unsigned idx = cur.next_idx++;
        cur.body.u16(idx); cur.body.u16(1); cur.body.u32(0); cur.body.u64(s.size());
        cur.body.raw(s.data(), s.size());
        cur.body.align8();
        elem.u32((uint32_t)s.size());
        elem.u64(cur.addr);
        elem.u32(idx);
    }

    // ---- object header messages ------------------------------------------------------------------------------
    static void msg(Bytes &oh, unsigned type, unsigned flags, const Bytes &body) {
        oh.u16(type);
        oh.u16((unsigned)pad8(body.size()));
        oh.u8(flags);
        oh.zeros(3);
        oh.raw(body.b.data(), body.size());
        oh.align8();
    }
    static Bytes dataspace_1d(uint64_t n, bool with_max = false) {
        Bytes m;
        m.u8(1); m.u8(1); m.u8(with_max ? 1 : 0); m.u8(0); m.u32(0);   // version 1, rank 1, flag 1 = max dims follow
        m.u64(n);
        if (with_max) m.u64(n);
        return m;
    }
    static Bytes datatype_u8() {
        Bytes m;
        m.u8(0x10); m.u8(0); m.u8(0); m.u8(0);          // version 1 class 0 (fixed point); little endian, unsigned
        m.u32(1);                                       // 1 byte
        m.u16(0); m.u16(8);                             // bit offset, precision
        return m;
    }
    static Bytes datatype_vlen_string() {
        Bytes m;
        m.u8(0x19); m.u8(0x01); m.u8(0); m.u8(0);       // version 1 class 9 (vlen); type = string, null-terminated, ASCII
        m.u32(16);                                      // length + heap address + index
        m.u8(0x10); m.u8(0); m.u8(0); m.u8(0);          // base type as the library encodes it for H5T_C_S1 strings:
        m.u32(1);                                       // a 1-byte fixed-point type
        m.u16(0); m.u16(8);
        return m;
    }
    void attribute(Bytes &oh, const Attr &a) {
        Bytes dt = datatype_vlen_string(), ds = dataspace_1d(1, true), m;
        m.u8(1); m.u8(0);
        m.u16((unsigned)a.name.size() + 1);
        m.u16((unsigned)dt.size());
        m.u16((unsigned)ds.size());
        m.raw(a.name.c_str(), a.name.size() + 1); m.align8();
        m.raw(dt.b.data(), dt.size()); m.align8();
        m.raw(ds.b.data(), ds.size()); m.align8();
        vlen_string(a.value, m);
        msg(oh, 0x000C, 0, m);
    }
    uint64_t object_header(unsigned nmsgs, const Bytes &msgs) {
        Bytes oh;
        oh.u8(1); oh.u8(0); oh.u16(nmsgs); oh.u32(1);
        oh.u32((uint32_t)msgs.size());
        oh.u32(0);                                      // pads the prefix to 16 bytes
        oh.raw(msgs.b.data(), msgs.size());
        uint64_t at = alloc(oh.size());
        put(at, std::move(oh));
        return at;
    }

    uint64_t write_dataset(const Dataset &d) {
        uint64_t data_at = H5_UNDEF;
        if (d.len) {
            data_at = alloc(d.len);
            put_ext(data_at, d.data, d.len);
        }
        Bytes msgs;
        msg(msgs, 0x0001, 0, dataspace_1d(d.len));
        msg(msgs, 0x0003, 1, datatype_u8());
        Bytes lay;
        lay.u8(3); lay.u8(1); lay.u64(data_at); lay.u64(d.len);    // version 3, contiguous
        msg(msgs, 0x0008, 0, lay);
        for (const Attr &a : d.attrs) attribute(msgs, a);
        return object_header(3 + (unsigned)d.attrs.size(), msgs);
    }

    // ---- groups ------------------------------------------------------------------------------------------------
    void write_group(const Group &g, uint64_t &ohdr, uint64_t &btree, uint64_t &heap) {
        std::vector<Link> links;
        for (const auto &sub : g.groups) {
            Link l;
            l.name = sub->name;
            l.is_group = true;
            write_group(*sub, l.ohdr, l.btree, l.heap);
            links.push_back(std::move(l));
        }
        for (const Dataset &d : g.dsets) {
            Link l;
            l.name = d.name;
            l.ohdr = write_dataset(d);
            links.push_back(std::move(l));
        }
        std::sort(links.begin(), links.end(), [](const Link &a, const Link &b) { return a.name < b.name; });

        // local heap: the empty name at offset 0, then every link name, each padded to 8 bytes
        Bytes names;
        names.zeros(8);
        for (Link &l : links) {
            l.name_off = names.size();
            names.raw(l.name.c_str(), l.name.size() + 1);
            names.align8();
        }
        Bytes hp;
        hp.raw("HEAP", 4); hp.u8(0); hp.zeros(3);
        hp.u64(names.size());
        hp.u64(1);                                      // no free block (H5HL_FREE_NULL)
        heap = alloc(32 + names.size());
        hp.u64(heap + 32);
        hp.raw(names.b.data(), names.size());
        put(heap, std::move(hp));

        // symbol nodes of up to 2K entries, then B-tree levels of up to 2K children until one root is left
        const size_t per_leaf = 2 * GROUP_LEAF_K, per_node = 2 * GROUP_INTERNAL_K;
        const uint64_t snod_size = 8 + per_leaf * 40, node_size = 24 + (2 * per_node + 1) * 8;
        struct Child { uint64_t addr; uint64_t last_key; };
        std::vector<Child> level;
        for (size_t i = 0; i < links.size(); i += per_leaf) {
            size_t n = std::min(per_leaf, links.size() - i);
            Bytes sn;
            sn.raw("SNOD", 4); sn.u8(1); sn.u8(0); sn.u16((unsigned)n);
            for (size_t j = i; j < i + n; j++) {
                const Link &l = links[j];
                sn.u64(l.name_off); sn.u64(l.ohdr);
                sn.u32(l.is_group ? 1 : 0); sn.u32(0);
                if (l.is_group) { sn.u64(l.btree); sn.u64(l.heap); } else sn.zeros(16);
            }
            sn.zeros((size_t)(snod_size - sn.size()));
            uint64_t at = alloc(snod_size);
            put(at, std::move(sn));
            level.push_back({at, links[i + n - 1].name_off});
        }
        unsigned depth = 0;
        for (;;) {
            size_t nnodes = std::max<size_t>(1, (level.size() + per_node - 1) / per_node);
            std::vector<uint64_t> addr(nnodes);
            for (size_t k = 0; k < nnodes; k++) addr[k] = alloc(node_size);
            std::vector<Child> up;
            for (size_t k = 0; k < nnodes; k++) {
                size_t lo = k * per_node, hi = std::min(level.size(), lo + per_node);
                Bytes nd;
                nd.raw("TREE", 4); nd.u8(0); nd.u8(depth); nd.u16((unsigned)(hi - lo));
                nd.u64(k ? addr[k - 1] : H5_UNDEF);
                nd.u64(k + 1 < nnodes ? addr[k + 1] : H5_UNDEF);
                nd.u64(lo ? level[lo - 1].last_key : 0);           // key 0: below every name of this subtree
                for (size_t c = lo; c < hi; c++) { nd.u64(level[c].addr); nd.u64(level[c].last_key); }
                nd.zeros((size_t)(node_size - nd.size()));
                put(addr[k], std::move(nd));
                up.push_back({addr[k], hi > lo ? level[hi - 1].last_key : 0});
            }
            level.swap(up);
            depth++;
            if (level.size() == 1) break;
        }
        btree = level[0].addr;

        Bytes msgs, st;
        st.u64(btree); st.u64(heap);
        msg(msgs, 0x0011, 0, st);
        ohdr = object_header(1, msgs);
    }

    int write_file(const char *path, const Group &root, std::string &err) {
        eof = 96;                                      // superblock v0 with 8-byte offsets and lengths
        uint64_t ohdr, btree, heap;
        write_group(root, ohdr, btree, heap);
        gcol_close();
        eof = pad8(eof);
        Bytes sb;
        sb.raw("\x89HDF\r\n\x1a\n", 8);
        sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0); sb.u8(0);   // superblock, free space, root entry, reserved, shared header
        sb.u8(8); sb.u8(8); sb.u8(0);
        sb.u16(GROUP_LEAF_K); sb.u16(GROUP_INTERNAL_K);
        sb.u32(0);
        sb.u64(0); sb.u64(H5_UNDEF); sb.u64(eof); sb.u64(H5_UNDEF);
        sb.u64(0); sb.u64(ohdr); sb.u32(1); sb.u32(0); sb.u64(btree); sb.u64(heap);
        put(0, std::move(sb));

        std::sort(pieces.begin(), pieces.end(), [](const Piece &a, const Piece &b) { return a.off < b.off; });
        FILE *f = fopen(path, "wb");
        if (!f) { err = std::string("cannot create ") + path + ": " + strerror(errno); return LEON_E_IO; }
        std::vector<char> iobuf(1 << 22);
        setvbuf(f, iobuf.data(), _IOFBF, iobuf.size());
        static const uint8_t zero[8] = {0};
        uint64_t pos = 0;
        bool ok = true;
        for (const Piece &p : pieces) {
            if (p.off < pos) { err = "container layout: overlapping extents"; ok = false; break; }
            while (ok && pos < p.off) {
                size_t n = (size_t)std::min<uint64_t>(8, p.off - pos);
                ok = fwrite(zero, 1, n, f) == n;
                pos += n;
            }
            const uint8_t *src = p.ext ? p.ext : p.own.data();
            if (ok && p.len) ok = fwrite(src, 1, p.len, f) == p.len;
            pos += p.len;
            if (!ok) break;
        }
        while (ok && pos < eof) { ok = fwrite(zero, 1, 1, f) == 1; pos++; }
        if (fclose(f) != 0) ok = false;
        if (!ok && err.empty()) err = std::string("write to ") + path + " failed: " + strerror(errno);
        return ok ? LEON_OK : LEON_E_IO;
    }
};

// zlib streams of the quality text of every block, as C/Leon.cpp:581-607 makes them (deflateInit at
// Z_BEST_COMPRESSION, one stream per block), blocks deflated by a pool of host threads.
int deflate_blocks(const uint8_t *text, const uint64_t *sizes, uint64_t nblocks, std::vector<std::vector<uint8_t>> &out,
                   std::string &err) {
    out.assign(nblocks, std::vector<uint8_t>());
    std::vector<uint64_t> start(nblocks + 1, 0);
    for (uint64_t b = 0; b < nblocks; b++) start[b + 1] = start[b] + sizes[b];
    std::atomic<uint64_t> next(0);
    std::atomic<int> failed(0);
    auto work = [&]() {
        for (;;) {
            uint64_t b = next.fetch_add(1);
            if (b >= nblocks || failed.load()) return;
            uLongf cap = compressBound((uLong)sizes[b]);
            out[b].resize(cap);
            int rc = compress2(out[b].data(), &cap, text + start[b], (uLong)sizes[b], Z_BEST_COMPRESSION);
            if (rc != Z_OK) { failed.store(rc ? rc : -1); return; }
            out[b].resize(cap);
        }
    };
    unsigned nthreads = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)std::min<uint64_t>(nblocks, 64)));
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nthreads; t++) pool.emplace_back(work);
    work();
    for (auto &t : pool) t.join();
    if (failed.load()) { err = "zlib compress2 failed on a quality block"; return LEON_E_IO; }
    return LEON_OK;
}

std::string dec(uint64_t v) { return std::to_string(v); }

}  // namespace

extern "C" int leon_write_container(const char *path, const leon_result *r, const leon_container_info *info) {
    std::string err;
    try {
        if (!path || !r || !info) { leon_internal_set_error("leon_write_container: null argument"); return LEON_E_ARG; }
        if (info->kmer_size < 2 || info->kmer_size > 63) { leon_internal_set_error("leon_write_container: bad kmer_size"); return LEON_E_ARG; }
        const bool fasta = info->is_fasta != 0, has_header = info->has_header != 0;
        const bool have_z = r->qual_z && r->qual_z_block_sizes;
        if (!fasta && r->nblocks && !have_z && (!r->qual_text || !r->qual_block_sizes)) {
            leon_internal_set_error("leon_write_container: FASTQ container needs the quality text or the zlib streams of the result");
            return LEON_E_ARG;
        }
        if (has_header && (!info->first_header || (info->header_nblocks && (!info->header_blocksizes || !info->header_data)))) {
            leon_internal_set_error("leon_write_container: has_header needs the header stream");
            return LEON_E_ARG;
        }

        Group root;
        // /metadata (C/Leon.cpp:500-531 infobyte + kmerSize, :1046-1070 read statistics)
        Group *meta = root.group("metadata");
        const uint8_t infobyte = (uint8_t)((fasta ? 1 : 0) | (has_header ? 0 : 2));
        const uint64_t ksize = (uint64_t)info->kmer_size, readcount = r->read_count, totaldna = r->total_dna_size;
        const int32_t minsz = r->min_read_size, maxsz = r->max_read_size;
        Dataset &ib = meta->dataset("infobyte", &infobyte, 1);
        ib.attrs.push_back({"type", fasta ? "fasta" : "fastq"});
        ib.attrs.push_back({"header", has_header ? "true" : "false"});
        ib.attrs.push_back({"version", "1.1.0"});
        meta->dataset("kmerSize", &ksize, 8);
        meta->dataset("readcount", &readcount, 8);
        meta->dataset("totalDnaSize", &totaldna, 8);
        meta->dataset("minSequenceSize", &minsz, 4);
        meta->dataset("maxSequenceSize", &maxsz, 4);

        // /leon/bloom (StorageTools.hpp:93-123)
        Group *leon = root.group("leon");
        Dataset &bl = leon->dataset("bloom", r->bloom, r->bloom_nbytes);
        bl.attrs.push_back({"size", dec(r->bloom_bits)});
        bl.attrs.push_back({"nb_hash", "7"});
        bl.attrs.push_back({"type", "neighbor"});
        bl.attrs.push_back({"kmer_size", dec(ksize)});

        // /leon/anchors (C/Leon.cpp:1234-1290)
        Group *anchors = leon->group("anchors");
        const uint64_t dict_size = r->anchors_dict_nbytes;
        const uint32_t anchor_count = (uint32_t)r->anchor_count;
        anchors->dataset("anchorsDict", r->anchors_dict, r->anchors_dict_nbytes);
        anchors->dataset("size", &dict_size, 8);
        anchors->dataset("anchorAdress", &anchor_count, 4);

        // /leon/dna (C/Leon.cpp:653-710 dna_<id> + "size", :1000-1020 nb_blocks = 2 * blocks, blocksizes)
        Group *dna = leon->group("dna");
        const uint64_t nb_entries = 2 * r->nblocks;
        uint64_t at = 0;
        for (uint64_t b = 0; b < r->nblocks; b++) {
            uint64_t n = r->blocksizes[2 * b];
            if (at + n > r->dna_nbytes) { leon_internal_set_error("leon_write_container: blocksizes exceed dna_nbytes"); return LEON_E_ARG; }
            if (n) {     // writeBlock() returns early on an empty block, C/Leon.cpp:654
                Dataset &d = dna->dataset("dna_" + dec(b), r->dna + at, n);
                d.attrs.push_back({"size", dec(n)});
            }
            at += n;
        }
        dna->dataset("nb_blocks", &nb_entries, 8);
        dna->dataset("blocksizes", r->blocksizes, nb_entries * 8);

        // /leon/qual (C/Leon.cpp:581-651), FASTQ only
        std::vector<std::vector<uint8_t>> zq;
        if (!fasta) {
            Group *qual = leon->group("qual");
            if (have_z) {                                   // streams written by the GPU (deflate.cuh)
                uint64_t zat = 0;
                for (uint64_t b = 0; b < r->nblocks; b++) {
                    const uint64_t n = r->qual_z_block_sizes[b];
                    if (zat + n > r->qual_z_nbytes) { leon_internal_set_error("leon_write_container: qual_z_block_sizes exceed qual_z_nbytes"); return LEON_E_ARG; }
                    Dataset &d = qual->dataset("qual_" + dec(b), r->qual_z + zat, n);
                    d.attrs.push_back({"size", dec(n)});
                    zat += n;
                }
            } else {
                int rc = deflate_blocks(r->qual_text, r->qual_block_sizes, r->nblocks, zq, err);
                if (rc) { leon_internal_set_error(err.c_str()); return rc; }
                for (uint64_t b = 0; b < r->nblocks; b++) {
                    Dataset &d = qual->dataset("qual_" + dec(b), zq[b].data(), zq[b].size());
                    d.attrs.push_back({"size", dec(zq[b].size())});
                }
            }
        }

        // /leon/header (C/Leon.cpp:714-800, :653-680): empty with -noheader / -seq-only
        Group *header = leon->group("header");
        uint64_t first_size = 0, header_entries = 0;
        if (has_header) {
            first_size = info->first_header_nbytes;
            header_entries = 2 * info->header_nblocks;
            header->dataset("firstheadersize", &first_size, 8);
            header->dataset("firstheader", info->first_header, first_size);
            header->dataset("nb_blocks", &header_entries, 8);
            header->dataset("blocksizes", info->header_blocksizes, header_entries * 8);
            uint64_t hat = 0;
            for (uint64_t b = 0; b < info->header_nblocks; b++) {
                uint64_t n = info->header_blocksizes[2 * b];
                if (n) {
                    Dataset &d = header->dataset("header_" + dec(b), info->header_data + hat, n);
                    d.attrs.push_back({"size", dec(n)});
                }
                hat += n;
            }
        }

        H5Writer w;
        int rc = w.write_file(path, root, err);
        if (rc) leon_internal_set_error(err.c_str());
        return rc;
    } catch (const std::bad_alloc &) {
        leon_internal_set_error("leon_write_container: out of host memory");
        return LEON_E_NOMEM;
    } catch (const std::exception &e) {
        leon_internal_set_error(e.what());
        return LEON_E_IO;
    }
}
