// `.leon` container reader (host only): the inverse of container.cpp for the decompression side.
// Reads both the files written here (contiguous datasets) and the ones the reference writes through the HDF5 library
// (tools/storage/impl/CollectionHDF5Patch.hpp:164-200: chunked, extendible, unfiltered u8 datasets), i.e. the subset
// of the HDF5 1.8 format those two writers produce: superblock v0, v1 object headers with continuation blocks,
// symbol-table groups (B-tree v1 + SNOD + local heap), dataspace v1/v2, layout v3 contiguous / chunked (B-tree v1 of
// raw-data chunks), v1 attribute messages with variable-length strings in global heap collections.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/leon_b200.h"

extern "C" void leon_internal_set_error(const char *msg);

namespace {

constexpr uint64_t UNDEF = ~0ull;

struct Bad : std::runtime_error {
    explicit Bad(const std::string &m) : std::runtime_error("not a readable .leon container: " + m) {}
};

struct ObjInfo {                       // one dataset, as far as this reader needs it
    uint64_t n = 0;                    // elements (bytes)
    int layout = -1;                   // 1 contiguous, 2 chunked
    uint64_t addr = UNDEF;             // data address or chunk B-tree address
    uint64_t chunk = 0;                // chunk length (chunked)
    std::map<std::string, std::string> attrs;
};

}  // namespace

struct leon_container {
    const uint8_t *p = nullptr;
    size_t n = 0;
    int fd = -1;
    uint64_t root_btree = 0, root_heap = 0;
    std::map<std::string, std::unique_ptr<std::vector<uint8_t>>> cache;   // assembled chunked datasets
    std::map<std::string, ObjInfo> infos;
    std::string attr_tmp;

    ~leon_container() {
        if (p && n) munmap((void *)p, n);
        if (fd >= 0) close(fd);
    }

    const uint8_t *at(uint64_t off, uint64_t len) const {
        if (off > n || len > n - off) throw Bad("address beyond the end of the file");
        return p + off;
    }
    uint64_t u64(uint64_t off) const { uint64_t v; memcpy(&v, at(off, 8), 8); return v; }
    uint32_t u32(uint64_t off) const { uint32_t v; memcpy(&v, at(off, 4), 4); return v; }
    uint16_t u16(uint64_t off) const { uint16_t v; memcpy(&v, at(off, 2), 2); return v; }
    uint8_t u8(uint64_t off) const { return *at(off, 1); }

    void open_superblock() {
        if (n < 96 || memcmp(p, "\x89HDF\r\n\x1a\n", 8) != 0) throw Bad("no HDF5 signature");
        if (u8(8) != 0 || u8(13) != 8 || u8(14) != 8) throw Bad("only superblock version 0 with 8-byte offsets is supported");
        if (u64(24) != 0) throw Bad("non-zero base address");
        // root symbol table entry at 56: name offset, object header, cache type, reserved, scratch (B-tree, heap)
        const uint64_t ohdr = u64(64);
        if (u32(72) == 1) { root_btree = u64(80); root_heap = u64(88); }
        else group_of(ohdr, &root_btree, &root_heap);
    }

    // ---- object headers ---------------------------------------------------------------------------------------
    struct Msg { unsigned type; uint64_t off; unsigned size; };
    std::vector<Msg> messages(uint64_t ohdr) const {
        if (u8(ohdr) != 1) throw Bad("only version 1 object headers are supported");
        unsigned nmsgs = u16(ohdr + 2);
        std::vector<Msg> out;
        std::vector<std::pair<uint64_t, uint64_t>> blocks{{ohdr + 16, u32(ohdr + 8)}};
        for (size_t b = 0; b < blocks.size() && out.size() < nmsgs; b++) {
            uint64_t q = blocks[b].first, end = q + blocks[b].second;
            while (q + 8 <= end && out.size() < nmsgs) {
                Msg m{u16(q), q + 8, u16(q + 2)};
                if (m.off + m.size > end) throw Bad("object header message overruns its block");
                if (m.type == 0x0010) blocks.push_back({u64(m.off), u64(m.off + 8)});   // continuation
                out.push_back(m);
                q = m.off + m.size;
            }
        }
        return out;
    }
    void group_of(uint64_t ohdr, uint64_t *btree, uint64_t *heap) const {
        for (const Msg &m : messages(ohdr))
            if (m.type == 0x0011) { *btree = u64(m.off); *heap = u64(m.off + 8); return; }
        throw Bad("group without a symbol table");
    }

    // ---- groups ---------------------------------------------------------------------------------------------
    struct Entry { uint64_t ohdr; bool is_group; uint64_t btree, heap; };
    bool find_in_tree(uint64_t node, uint64_t heap_data, const std::string &name, Entry *out, int depth = 0) const {
        if (depth > 32) throw Bad("group B-tree too deep");
        if (memcmp(at(node, 4), "TREE", 4) == 0) {
            if (u8(node + 4) != 0) throw Bad("unexpected B-tree type in a group");
            const unsigned used = u16(node + 6);
            for (unsigned i = 0; i < used; i++)
                if (find_in_tree(u64(node + 24 + 8 + 16ull * i), heap_data, name, out, depth + 1)) return true;
            return false;
        }
        if (memcmp(at(node, 4), "SNOD", 4) != 0) throw Bad("bad symbol table node");
        const unsigned nsym = u16(node + 6);
        for (unsigned i = 0; i < nsym; i++) {
            const uint64_t e = node + 8 + 40ull * i;
            const char *nm = (const char *)at(heap_data + u64(e), 1);
            const size_t maxlen = n - (size_t)(heap_data + u64(e));
            if (strnlen(nm, maxlen) == name.size() && memcmp(nm, name.data(), name.size()) == 0) {
                out->ohdr = u64(e + 8);
                out->is_group = false;
                if (u32(e + 16) == 1) { out->is_group = true; out->btree = u64(e + 24); out->heap = u64(e + 32); }
                else {
                    for (const Msg &m : messages(out->ohdr))
                        if (m.type == 0x0011) { out->is_group = true; out->btree = u64(m.off); out->heap = u64(m.off + 8); }
                }
                return true;
            }
        }
        return false;
    }
    bool lookup(const std::string &path, Entry *out) const {
        uint64_t btree = root_btree, heap = root_heap;
        size_t i = 0;
        Entry e{0, true, btree, heap};
        while (i < path.size()) {
            while (i < path.size() && path[i] == '/') i++;
            size_t j = path.find('/', i);
            if (j == std::string::npos) j = path.size();
            if (j == i) break;
            if (!e.is_group) return false;
            if (memcmp(at(e.heap, 4), "HEAP", 4) != 0) throw Bad("bad local heap");
            const uint64_t heap_data = u64(e.heap + 24);
            Entry nxt;
            if (!find_in_tree(e.btree, heap_data, path.substr(i, j - i), &nxt)) return false;
            e = nxt;
            i = j;
        }
        *out = e;
        return true;
    }

    // ---- datasets and attributes ------------------------------------------------------------------------------
    std::string vlen_string(uint64_t elem) const {
        const uint32_t len = u32(elem);
        const uint64_t col = u64(elem + 4);
        const uint32_t idx = u32(elem + 12);
        if (!len || !col) return std::string();
        if (memcmp(at(col, 4), "GCOL", 4) != 0) throw Bad("bad global heap collection");
        const uint64_t size = u64(col + 8);
        uint64_t q = col + 16;
        while (q + 16 <= col + size) {
            const unsigned i = u16(q);
            const uint64_t osz = u64(q + 8);
            if (i == 0) break;
            if (i == idx) return std::string((const char *)at(q + 16, osz), (size_t)std::min<uint64_t>(osz, len));
            q += 16 + ((osz + 7) & ~7ull);
        }
        throw Bad("global heap object not found");
    }

    const ObjInfo &info(const std::string &path) {
        auto it = infos.find(path);
        if (it != infos.end()) return it->second;
        Entry e;
        if (!lookup(path, &e) || e.is_group) throw Bad("dataset " + path + " not found");
        ObjInfo o;
        for (const Msg &m : messages(e.ohdr)) {
            if (m.type == 0x0001) {                                 // dataspace
                const unsigned ver = u8(m.off), rank = u8(m.off + 1);
                if (rank != 1) throw Bad("dataset " + path + " is not one-dimensional");
                o.n = u64(m.off + (ver == 1 ? 8 : 4));
            } else if (m.type == 0x0003) {                          // datatype: 1-byte fixed point expected
                if ((u8(m.off) & 15) != 0 || u32(m.off + 4) != 1) throw Bad("dataset " + path + " is not a byte array");
            } else if (m.type == 0x0008) {                          // layout
                if (u8(m.off) != 3) throw Bad("only version 3 data layouts are supported");
                o.layout = u8(m.off + 1);
                if (o.layout == 1) o.addr = u64(m.off + 2);
                else if (o.layout == 2) {
                    if (u8(m.off + 2) != 2) throw Bad("chunked dataset " + path + " is not one-dimensional");
                    o.addr = u64(m.off + 3);
                    o.chunk = u32(m.off + 11);
                } else throw Bad("unsupported layout class in " + path);
            } else if (m.type == 0x000B) {
                throw Bad("dataset " + path + " uses a filter pipeline");
            } else if (m.type == 0x000C) {                          // attribute, version 1
                if (u8(m.off) != 1) throw Bad("only version 1 attribute messages are supported");
                const unsigned nl = u16(m.off + 2), dl = u16(m.off + 4), sl = u16(m.off + 6);
                uint64_t q = m.off + 8;
                std::string name((const char *)at(q, nl), nl ? nl - 1 : 0);
                q += (nl + 7) & ~7u;
                const bool vlen_str = (u8(q) & 15) == 9 && (u8(q + 1) & 15) == 1;
                const uint32_t fixed = u32(q + 4);
                const bool fixed_str = (u8(q) & 15) == 3;
                q += (dl + 7) & ~7u;
                q += (sl + 7) & ~7u;
                if (vlen_str) o.attrs[name] = vlen_string(q);
                else if (fixed_str) o.attrs[name] = std::string((const char *)at(q, fixed), strnlen((const char *)at(q, fixed), fixed));
            }
        }
        if (o.layout < 0) throw Bad("dataset " + path + " has no layout");
        return infos.emplace(path, std::move(o)).first->second;
    }

    void gather_chunks(uint64_t node, const ObjInfo &o, std::vector<uint8_t> &dst, int depth = 0) const {
        if (depth > 32) throw Bad("chunk B-tree too deep");
        if (memcmp(at(node, 4), "TREE", 4) != 0 || u8(node + 4) != 1) throw Bad("bad chunk B-tree node");
        const unsigned level = u8(node + 5), used = u16(node + 6);
        const uint64_t keysz = 4 + 4 + 8 * 2;                      // chunk bytes, filter mask, offsets of rank + 1 dimensions
        for (unsigned i = 0; i < used; i++) {
            const uint64_t key = node + 24 + (keysz + 8) * i;
            const uint64_t child = u64(key + keysz);
            if (level) { gather_chunks(child, o, dst, depth + 1); continue; }
            const uint32_t nbytes = u32(key);
            if (u32(key + 4) != 0) throw Bad("filtered chunk");
            const uint64_t off = u64(key + 8);
            if (off >= o.n) continue;
            const uint64_t take = std::min<uint64_t>(std::min<uint64_t>(nbytes, o.chunk), o.n - off);
            memcpy(dst.data() + off, at(child, take), take);
        }
    }

    const uint8_t *dataset(const std::string &path, uint64_t *len) {
        const ObjInfo &o = info(path);
        *len = o.n;
        if (o.n == 0) return p;
        if (o.layout == 1) {
            if (o.addr == UNDEF) throw Bad("dataset " + path + " has no storage");
            return at(o.addr, o.n);
        }
        auto it = cache.find(path);
        if (it == cache.end()) {
            std::unique_ptr<std::vector<uint8_t>> v(new std::vector<uint8_t>(o.n, 0));
            if (o.addr != UNDEF) gather_chunks(o.addr, o, *v);
            it = cache.emplace(path, std::move(v)).first;
        }
        return it->second->data();
    }
};

extern "C" int leon_container_open(const char *path, leon_container **out) {
    try {
        if (!path || !out) { leon_internal_set_error("leon_container_open: null argument"); return LEON_E_ARG; }
        std::unique_ptr<leon_container> c(new leon_container());
        c->fd = open(path, O_RDONLY);
        struct stat st;
        if (c->fd < 0 || fstat(c->fd, &st) != 0) {
            leon_internal_set_error((std::string("cannot open ") + path + ": " + strerror(errno)).c_str());
            return LEON_E_IO;
        }
        c->n = (size_t)st.st_size;
        if (c->n) {
            void *q = mmap(nullptr, c->n, PROT_READ, MAP_PRIVATE, c->fd, 0);
            if (q == MAP_FAILED) { c->n = 0; leon_internal_set_error("cannot map the container"); return LEON_E_IO; }
            c->p = (const uint8_t *)q;
        }
        c->open_superblock();
        *out = c.release();
        return LEON_OK;
    } catch (const std::exception &e) {
        leon_internal_set_error(e.what());
        return LEON_E_IO;
    }
}

extern "C" void leon_container_close(leon_container *c) { delete c; }

extern "C" int leon_container_dataset(leon_container *c, const char *path, const uint8_t **data, uint64_t *nbytes) {
    try {
        if (!c || !path || !data || !nbytes) { leon_internal_set_error("leon_container_dataset: null argument"); return LEON_E_ARG; }
        *data = c->dataset(path, nbytes);
        return LEON_OK;
    } catch (const std::exception &e) {
        leon_internal_set_error(e.what());
        return LEON_E_IO;
    }
}

extern "C" int leon_container_has(leon_container *c, const char *path) {
    try {
        leon_container::Entry e;
        return c && path && c->lookup(path, &e) ? 1 : 0;
    } catch (const std::exception &) {
        return 0;
    }
}

extern "C" int leon_container_attr(leon_container *c, const char *dataset_path, const char *name, const char **value) {
    try {
        if (!c || !dataset_path || !name || !value) { leon_internal_set_error("leon_container_attr: null argument"); return LEON_E_ARG; }
        const ObjInfo &o = c->info(dataset_path);
        auto it = o.attrs.find(name);
        if (it == o.attrs.end()) {
            leon_internal_set_error((std::string("attribute ") + name + " of " + dataset_path + " not found").c_str());
            return LEON_E_IO;
        }
        *value = it->second.c_str();
        return LEON_OK;
    } catch (const std::exception &e) {
        leon_internal_set_error(e.what());
        return LEON_E_IO;
    }
}
