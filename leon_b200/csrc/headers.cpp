// Header stream of `leon -c` on the host: the read names are split into fields and coded
// against the previous header with adaptive order-0 models and the 64-bit range coder, one independent stream per
// block of reads_per_block reads.  Follows C/HeaderCoder.cpp:33-213 (field splitting, per-column models,
// startBlock), :257-524 (HeaderEncoder: compareHeader, encodeNumeric, encodeAscii), C/CompressionUtils.hpp:75-103
// (varint numerics) and :162-199 (delta choice), C/RangeCoder.cpp:29-145 (Order0Model, RangeEncoder), and
// C/Leon.cpp:786-899 (first header = first comment minus its first character, block table).
//
// The reference keeps ONE encoder object per thread for the whole file and never resets some of its members between
// headers or blocks (slots of the field arrays past the current field count, _lastMatchFieldIndex, numeric models 8..19):
// with -nb-cores 1 the stream therefore depends, in rare cases, on state left by earlier blocks.  This coder codes the
// blocks in parallel with a fresh object per block, records whether any such left-over state was consulted, and in that
// case recodes the whole file with one serial object exactly like the reference with -nb-cores 1; either way the
// result is byte-identical to it.
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/leon_b200.h"
#include "decoded.h"

extern "C" void leon_internal_set_error(const char *msg);

namespace {

enum HeaderType { HEADER_END = 1, HEADER_END_MATCH, FIELD_ASCII, FIELD_NUMERIC, FIELD_DELTA, FIELD_DELTA_2, FIELD_ZERO_ONLY,
                  FIELD_ZERO_AND_NUMERIC, HEADER_TYPE_COUNT };
constexpr int NB_MODELS_PER_NUMERIC = 20;
constexpr uint64_t RC_TOP = 1ull << 56, RC_BOTTOM = 1ull << 48;

// Adaptive order-0 model, every symbol starting at frequency 1 and +1 per occurrence (C/RangeCoder.cpp:29-81).  The
// reference keeps the cumulative table and adds 1 to every entry above the symbol; here counts are kept per symbol
// with sums over groups of 16, which gives the same low / high / total in ~32 additions.  The halving at a total of
// 2^48 cannot be reached by a block (it would need 2^48 symbols); counts are 32 bit.
struct Model {
    int n = 0;
    std::vector<uint32_t> cnt, grp;
    explicit Model(int symbols = 0) { init(symbols); }
    void init(int symbols) { n = symbols; cnt.assign((size_t)((n + 15) & ~15), 0); grp.assign(cnt.size() / 16, 0); clear(); }
    void clear() {
        for (size_t i = 0; i < cnt.size(); i++) cnt[i] = i < (size_t)n ? 1 : 0;
        for (size_t g = 0; g < grp.size(); g++) { uint32_t s = 0; for (int j = 0; j < 16; j++) s += cnt[g * 16 + j]; grp[g] = s; }
        total = (uint64_t)n;
    }
    uint64_t total = 0;
    uint64_t low(unsigned c) const {
        uint64_t s = 0;
        for (unsigned g = 0; g < (c >> 4); g++) s += grp[g];
        for (unsigned i = c & ~15u; i < c; i++) s += cnt[i];
        return s;
    }
    void update(unsigned c) { cnt[c]++; grp[c >> 4]++; total++; }
    // largest symbol c with low(c) <= count (RangeDecoder::nextByte, C/RangeCoder.cpp:194-206); *lo gets low(c)
    unsigned find(uint64_t count, uint64_t *lo) const {
        uint64_t s = 0;
        size_t g = 0;
        while (g + 1 < grp.size() && s + grp[g] <= count) { s += grp[g]; g++; }
        unsigned c = (unsigned)g * 16;
        while ((int)c < n - 1 && s + cnt[c] <= count) { s += cnt[c]; c++; }
        *lo = s;
        return c;
    }
};

struct RangeEncoder {
    uint64_t low = 0, range = ~0ull;
    std::vector<uint8_t> buf;
    void encode(Model &m, unsigned c) {                  // C/RangeCoder.cpp:106-134
        range /= m.total;
        low += m.low(c) * range;
        range *= m.cnt[c];
        while ((low ^ (low + range)) < RC_TOP || range < RC_BOTTOM) {
            if (range < RC_BOTTOM && (low ^ (low + range)) >= RC_TOP) range = (0 - low) & (RC_BOTTOM - 1);
            buf.push_back((uint8_t)(low >> 56));
            range <<= 8;
            low <<= 8;
        }
        m.update(c);
    }
    void flush() {                                       // C/RangeCoder.cpp:136-145
        for (int i = 0; i < 8; i++) { buf.push_back((uint8_t)(low >> 56)); low <<= 8; }
        low = 0; range = ~0ull;
    }
    void clear() { low = 0; range = ~0ull; buf.clear(); }
};

struct RangeDecoder {                                    // C/RangeCoder.cpp:179-244
    uint64_t low = 0, range = ~0ull, code = 0;
    const uint8_t *p = nullptr, *end = nullptr;
    uint8_t next() { return p < end ? *p++ : (uint8_t)0xff; }
    void start(const uint8_t *data, uint64_t n) {
        low = 0; range = ~0ull; code = 0; p = data; end = data + n;
        for (int i = 0; i < 8; i++) code = (code << 8) | next();
    }
    unsigned decode(Model &m) {
        range /= m.total;
        uint64_t lo;
        const unsigned c = m.find((code - low) / range, &lo);
        low += lo * range;
        range *= m.cnt[c];
        while ((low ^ (low + range)) < RC_TOP || range < RC_BOTTOM) {
            if (range < RC_BOTTOM && (low ^ (low + range)) >= RC_TOP) range = (0 - low) & (RC_BOTTOM - 1);
            code = (code << 8) | next();
            range <<= 8;
            low <<= 8;
        }
        m.update(c);
        return c;
    }
};

struct Column {                                          // AbstractHeaderCoder::addFieldColumn, C/HeaderCoder.cpp:41-62
    Model type{HEADER_TYPE_COUNT + 1}, field_index{256}, field_column{256}, mis_size{256}, ascii{128}, zero{256};
    std::vector<Model> numeric;
    unsigned prev_pos = 0, cur_pos = 0;
    uint64_t prev_value = 0, cur_value = 0, prev_zero = 0, cur_zero = 0;
    int prev_type = FIELD_ASCII, cur_type = FIELD_ASCII;
    Column() { numeric.assign(NB_MODELS_PER_NUMERIC, Model(256)); }
};

struct View { const char *p; size_t n; char at(size_t i) const { return i < n ? p[i] : '\0'; } };

// AbstractHeaderCoder (C/HeaderCoder.cpp:33-213): field splitting and the per-column state both directions share.
class HeaderCoderBase {
  public:
    bool leftover_used = false;     // a member the reference never resets was consulted before this object set it
    bool bad_char = false;          // a character >= 128 (outside the reference's 128-symbol text model)

    explicit HeaderCoderBase(View first) : first_(first) {}

    void start_block() {                                 // C/HeaderCoder.cpp:188-213
        cur_ = first_;
        for (Column &c : col_) {
            c.type.clear(); c.field_index.clear(); c.field_column.clear(); c.mis_size.clear(); c.ascii.clear(); c.zero.clear();
            for (int j = 0; j < 8; j++) c.numeric[j].clear();         // the reference clears only 8 of the 20
        }
        header_size_.clear();
        split_header();
        end_header();
        processed_ = 0;
    }
    int processed() const { return processed_; }

  protected:
    View first_, prev_{nullptr, 0}, cur_{nullptr, 0};
    std::vector<Column> col_;
    Model header_size_{256};
    int field_index_ = 0, field_start_ = 0, pos_ = 0, cur_count_ = 0, prev_count_ = 0, mis_index_ = 0, processed_ = 0;
    int last_match_ = 0;
    bool last_match_set_ = false, numeric_ = true;
    int cur_size_ = 0, prev_size_ = 0, mis_start_ = -1, field_pos_ = 0;

    static int type_of_char(unsigned char c, bool *digit) {          // C/HeaderCoder.cpp:64-77 ("C" locale)
        if (c >= '0' && c <= '9') { *digit = true; return 1; }
        *digit = false;
        return ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) ? 1 : 2;
    }
    void add_column() { col_.emplace_back(); }

    void split_header() {                                // C/HeaderCoder.cpp:79-109
        field_index_ = 0; field_start_ = 0; pos_ = 0; numeric_ = true;
        bool digit_only;
        int last_type = type_of_char((unsigned char)cur_.at(0), &digit_only);
        for (pos_ = 0; (size_t)pos_ < cur_.n; pos_++) {
            unsigned char c = (unsigned char)cur_.p[pos_];
            if (c >= 128) bad_char = true;
            int t = type_of_char(c, &digit_only);
            if (t != last_type) { last_type = t; make_field(); }
            if (numeric_) numeric_ = digit_only;
        }
        make_field();
        cur_count_ = field_index_;
    }
    void make_field() {                                  // C/HeaderCoder.cpp:111-164
        if (field_start_ == pos_) return;
        while ((int)col_.size() <= field_index_ + 1) add_column();
        Column &c = col_[field_index_];
        c.cur_pos = (unsigned)field_start_;
        col_[field_index_ + 1].cur_pos = (unsigned)pos_;
        if (numeric_) {
            const char *f = cur_.p + field_start_;
            int len = pos_ - field_start_, zeros = 0;
            while (zeros < len && f[zeros] == '0') zeros++;
            c.cur_zero = (uint64_t)zeros;
            // strtoul(digits, NULL, 0) on a string without leading zeros: decimal, saturating at ULONG_MAX
            uint64_t v = 0;
            bool sat = false;
            for (int i = zeros; i < len; i++) {
                unsigned d = (unsigned)(f[i] - '0');
                if (v > (~0ull - d) / 10) { sat = true; break; }
                v = v * 10 + d;
            }
            c.cur_value = sat ? ~0ull : v;
            c.cur_type = zeros > 0 ? (c.cur_value == 0 ? FIELD_ZERO_ONLY : FIELD_ZERO_AND_NUMERIC) : FIELD_NUMERIC;
        } else {
            c.cur_type = FIELD_ASCII;
        }
        field_index_ += 1;
        field_start_ = pos_;
        numeric_ = true;
    }
    void end_header() {                                  // C/HeaderCoder.cpp:166-186
        prev_count_ = cur_count_;
        for (int i = 0; i < prev_count_ + 1 && i < (int)col_.size(); i++) {
            Column &c = col_[i];
            c.prev_pos = c.cur_pos; c.prev_value = c.cur_value; c.prev_type = c.cur_type; c.prev_zero = c.cur_zero;
            c.cur_zero = 0;
        }
        prev_ = cur_;
        mis_index_ = 0;
        field_index_ = 0;
        processed_ += 1;
    }

    Column &col_at(int i) {
        while ((int)col_.size() <= i) add_column();      // the reference indexes its vectors unchecked; only reachable
        return col_[i];                                  // for an empty header on a fresh object
    }
};

class HeaderEncoder : public HeaderCoderBase {
  public:
    RangeEncoder rc;
    explicit HeaderEncoder(View first) : HeaderCoderBase(first) {}
    void encode(View h) {                                // processNextHeader, C/HeaderCoder.cpp:290-299
        cur_ = h;
        split_header();
        compare_header();
        end_header();
    }

  private:
    void compare_header() {                              // C/HeaderCoder.cpp:301-426
        field_pos_ = 0;
        mis_start_ = -1;
        for (field_index_ = 0; field_index_ < cur_count_; field_index_++) {
            Column &c = col_[field_index_];
            cur_size_ = (int)(col_[field_index_ + 1].cur_pos - c.cur_pos);
            prev_size_ = (int)(col_[field_index_ + 1].prev_pos - c.prev_pos);
            mis_start_ = -1;
            const int pt = c.prev_type, ct = c.cur_type;
            const bool numeric_like = ct == FIELD_NUMERIC || ct == FIELD_ZERO_ONLY || ct == FIELD_ZERO_AND_NUMERIC;
            // Past the previous header's fields the "previous" slots hold whatever an earlier header left there.
            if (numeric_like && field_index_ >= prev_count_) leftover_used = true;
            if (pt == FIELD_NUMERIC && ct == FIELD_NUMERIC) {
                if (c.prev_value == c.cur_value) { set_match(); continue; }
            } else if (pt == FIELD_ZERO_ONLY && ct == FIELD_ZERO_ONLY) {
                if (c.prev_zero == c.cur_zero) { set_match(); continue; }
            } else if (pt == FIELD_ZERO_AND_NUMERIC && ct == FIELD_ZERO_AND_NUMERIC) {
                if (c.prev_zero == c.cur_zero && c.prev_value == c.cur_value) { set_match(); continue; }
            }
            if (numeric_like) {
                encode_numeric();
            } else {
                for (field_pos_ = 0; field_pos_ < cur_size_; field_pos_++) {
                    if (field_index_ >= prev_count_) { mis_start_ = field_pos_; break; }
                    if (field_pos_ >= prev_size_) { mis_start_ = field_pos_; break; }
                    if (cur_.p[c.cur_pos + field_pos_] != prev_.p[c.prev_pos + field_pos_]) { mis_start_ = field_pos_; break; }
                }
                if (mis_start_ != -1) {
                    encode_ascii();
                } else if (field_pos_ != prev_size_) {
                    mis_start_ = field_pos_;
                    encode_ascii();
                } else {
                    set_match();
                }
            }
        }
        if (!last_match_set_) leftover_used = true;      // _lastMatchFieldIndex is read here whether or not it was set
        if (last_match_set_ && last_match_ == field_index_ - 1) {
            rc.encode(col_at(mis_index_).type, HEADER_END_MATCH);
            rc.encode(header_size_, (unsigned)(cur_.n & 255));
        } else {
            rc.encode(col_at(mis_index_).type, HEADER_END);
        }
    }
    void set_match() { last_match_ = field_index_; last_match_set_ = true; }

    void encode_numeric() {                              // C/HeaderCoder.cpp:428-489
        Column &f = col_[field_index_];
        const uint64_t zeros = f.cur_zero, field_value = f.cur_value;
        if (f.cur_type == FIELD_ZERO_ONLY) {
            Column &m = col_at(mis_index_);
            rc.encode(m.type, FIELD_ZERO_ONLY);
            rc.encode(m.field_index, (unsigned)(field_index_ & 255));
            rc.encode(m.zero, (unsigned)(zeros & 255));
            mis_index_ += 1;
            return;
        }
        if (f.cur_type == FIELD_ZERO_AND_NUMERIC) {
            Column &m = col_at(mis_index_);
            rc.encode(m.type, FIELD_ZERO_AND_NUMERIC);
            rc.encode(m.field_index, (unsigned)(field_index_ & 255));
            rc.encode(m.zero, (unsigned)(zeros & 255));
            mis_index_ += 1;
        }
        uint64_t value = field_value;
        const uint64_t prev = f.prev_value;
        // CompressionUtils::getDeltaValue, C/CompressionUtils.hpp:162-199
        const uint64_t d1 = value - prev, d2 = prev - value;
        const bool v1 = d1 < value, v2 = d2 < value;
        int delta_type = 0;
        if (v1 && v2) { if (d1 <= d2) { delta_type = 1; value = d1; } else { delta_type = 2; value = d2; } }
        else if (v1) { delta_type = 1; value = d1; }
        else if (v2) { delta_type = 2; value = d2; }
        Column &m = col_at(mis_index_);
        rc.encode(m.type, delta_type == 0 ? FIELD_NUMERIC : delta_type == 1 ? FIELD_DELTA : FIELD_DELTA_2);
        rc.encode(m.field_index, (unsigned)(field_index_ & 255));
        int i = 0;                                       // CompressionUtils::encodeNumeric, C/CompressionUtils.hpp:75-103
        do {
            unsigned byte = (unsigned)(value & 127);
            value >>= 7;
            if (value != 0) byte |= 128;
            if (i >= 8) leftover_used = true;            // models 8..19 are not cleared at block starts
            rc.encode(m.numeric[i], byte);
            i++;
        } while (value != 0);
        mis_index_ += 1;
    }
    void encode_ascii() {                                // C/HeaderCoder.cpp:491-524
        Column &f = col_[field_index_];
        Column &m = col_at(mis_index_);
        const int miss = cur_size_ - mis_start_;
        rc.encode(m.type, FIELD_ASCII);
        rc.encode(m.field_index, (unsigned)(field_index_ & 255));
        rc.encode(m.field_column, (unsigned)(mis_start_ & 255));
        rc.encode(m.mis_size, (unsigned)(miss & 255));
        for (int i = mis_start_; i < mis_start_ + miss; i++) rc.encode(m.ascii, (unsigned char)cur_.p[f.cur_pos + i] & 127);
        mis_index_ += 1;
    }
};

// HeaderDecoder (C/HeaderCoder.cpp:526-789): rebuilds every header from the previous one and the coded differences.
class HeaderDecoder : public HeaderCoderBase {
  public:
    RangeDecoder rc;
    explicit HeaderDecoder(View first) : HeaderCoderBase(first) {}

    // Decodes `count` headers of one header_<id> block into out, one per line (HeaderDecoder::_buffer).
    void run(const uint8_t *data, uint64_t nbytes, uint64_t count, std::string &out) {
        start_block();
        rc.start(data, nbytes);
        buf_.clear();
        mis_index_ = 0;
        field_index_ = 0;
        while ((uint64_t)processed_ < count) {
            const unsigned type = rc.decode(col_at(mis_index_).type);
            if (type == HEADER_END) {
                finish_header(out);
            } else if (type == HEADER_END_MATCH) {
                const unsigned header_size = rc.decode(header_size_);     // 8 bits of the real size, as coded
                for (; field_index_ < prev_count_; field_index_++) {
                    append_prev_field(field_index_);
                    if (buf_.size() >= header_size) break;
                }
                finish_header(out);
            } else {
                const int mis_field = (int)rc.decode(col_at(mis_index_).field_index);    // decodeMatch
                for (; field_index_ < mis_field; field_index_++) append_prev_field(field_index_);
                if (type == FIELD_ASCII) {
                    Column &m = col_at(mis_index_);
                    const unsigned column = rc.decode(m.field_column), size = rc.decode(m.mis_size);
                    if (field_index_ < prev_count_)
                        for (unsigned i = 0; i < column; i++) buf_ += prev_.at(col_at(field_index_).prev_pos + i);
                    for (unsigned i = 0; i < size; i++) buf_ += (char)rc.decode(m.ascii);
                    field_index_++; mis_index_++;
                } else if (type == FIELD_NUMERIC || type == FIELD_DELTA || type == FIELD_DELTA_2) {
                    uint64_t v = numeric(col_at(mis_index_).numeric);
                    if (type != FIELD_NUMERIC) {
                        if (field_index_ >= prev_count_) leftover_used = true;       // previous value of a field the previous header lacks
                        const uint64_t prev = col_at(field_index_).prev_value;
                        v = type == FIELD_DELTA ? prev + v : prev - v;
                    }
                    buf_ += std::to_string(v);
                    field_index_++; mis_index_++;
                } else if (type == FIELD_ZERO_ONLY || type == FIELD_ZERO_AND_NUMERIC) {
                    const unsigned zeros = rc.decode(col_at(mis_index_).zero);
                    buf_.append(zeros, '0');
                    if (type == FIELD_ZERO_ONLY) field_index_++;
                    mis_index_++;
                } else {
                    bad_stream = true;                                               // not a header type: corrupt block
                    return;
                }
            }
            if (buf_.size() > (1u << 20)) { bad_stream = true; return; }
        }
    }
    bool bad_stream = false;

  private:
    std::string buf_, prev_buf_;
    void append_prev_field(int f) {
        if (f + 1 >= (int)col_.size()) return;
        const unsigned a = col_[f].prev_pos, b = col_[f + 1].prev_pos;
        if (a > prev_.n) { bad_stream = true; return; }                              // the reference's substr throws here
        if (b > a) buf_.append(prev_.p + a, std::min<size_t>(b - a, prev_.n - a));
    }
    uint64_t numeric(std::vector<Model> &models) {                                   // C/CompressionUtils.hpp:114-137
        uint64_t value = 0;
        int shift = 0, i = 0;
        unsigned byte;
        do {
            if (i >= 8) leftover_used = true;
            byte = rc.decode(models[i < NB_MODELS_PER_NUMERIC ? i : NB_MODELS_PER_NUMERIC - 1]);
            if (shift < 64) value += (uint64_t)(byte & 127) << shift;
            shift += 7;
            i++;
        } while (byte > 127);
        return value;
    }
    void finish_header(std::string &out) {                                           // HeaderDecoder::endHeader
        out += buf_;
        out += '\n';
        cur_ = View{buf_.data(), buf_.size()};
        split_header();
        end_header();                                                                // prev_ = cur_ (a view of buf_)
        prev_buf_.swap(buf_);
        prev_ = View{prev_buf_.data(), prev_buf_.size()};
        buf_.clear();
        mis_index_ = 0;
    }
};

struct Stream {
    leon_header_stream pub;
    std::string first;
    std::vector<uint64_t> blocksizes;
    std::vector<uint8_t> data;
};

}  // namespace

extern "C" int leon_headers_encode(const uint8_t *text, const uint64_t *offsets, uint64_t nreads, int32_t reads_per_block,
                                   int32_t nthreads, leon_header_stream **out) {
    try {
        if (!out || (nreads && (!text || !offsets)) || reads_per_block <= 0) {
            leon_internal_set_error("leon_headers_encode: bad argument");
            return LEON_E_ARG;
        }
        std::unique_ptr<Stream> s(new Stream());
        const char *t = (const char *)text;
        auto header = [&](uint64_t i) { return View{t + offsets[i], (size_t)(offsets[i + 1] - offsets[i])}; };
        if (nreads && offsets[1] > offsets[0]) s->first.assign(t + offsets[0] + 1, (size_t)(offsets[1] - offsets[0] - 1));
        const View first{s->first.data(), s->first.size()};
        const uint64_t rpb = (uint64_t)reads_per_block, nblocks = (nreads + rpb - 1) / rpb;
        std::vector<std::vector<uint8_t>> blocks(nblocks);
        std::atomic<int> leftover(0), bad(0);

        auto code_block = [&](HeaderEncoder &e, uint64_t b) {
            e.start_block();
            const uint64_t lo = b * rpb, hi = std::min(nreads, lo + rpb);
            for (uint64_t i = lo; i < hi; i++) e.encode(header(i));
            if (!e.rc.buf.empty()) e.rc.flush();          // HeaderEncoder::writeBlock, C/HeaderCoder.cpp:277-288
            blocks[b].swap(e.rc.buf);
            e.rc.clear();
            if (e.bad_char) bad.store(1);
        };

        unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::max(1u, std::thread::hardware_concurrency());
        nt = (unsigned)std::min<uint64_t>(nt, std::max<uint64_t>(nblocks, 1));
        bool serial = nt <= 1;
        if (!serial) {
            std::atomic<uint64_t> next(0);
            auto work = [&]() {
                for (;;) {
                    uint64_t b = next.fetch_add(1);
                    if (b >= nblocks || leftover.load()) return;
                    HeaderEncoder e(first);
                    code_block(e, b);
                    if (e.leftover_used) leftover.store(1);
                }
            };
            std::vector<std::thread> pool;
            for (unsigned i = 1; i < nt; i++) pool.emplace_back(work);
            work();
            for (auto &th : pool) th.join();
            serial = leftover.load() != 0;
        }
        if (serial) {                                    // one object for the whole file = the reference with -nb-cores 1
            HeaderEncoder e(first);
            for (uint64_t b = 0; b < nblocks; b++) code_block(e, b);
        }
        if (bad.load()) {
            leon_internal_set_error("leon_headers_encode: a header holds a byte >= 128, outside the reference's 128-symbol text model");
            return LEON_E_ARG;
        }
        s->blocksizes.resize(2 * nblocks);
        uint64_t total = 0;
        for (uint64_t b = 0; b < nblocks; b++) total += blocks[b].size();
        s->data.reserve(total);
        for (uint64_t b = 0; b < nblocks; b++) {
            s->blocksizes[2 * b] = blocks[b].size();
            s->blocksizes[2 * b + 1] = std::min(nreads, (b + 1) * rpb) - b * rpb;
            s->data.insert(s->data.end(), blocks[b].begin(), blocks[b].end());
        }
        s->pub.first_header = (const uint8_t *)s->first.data();
        s->pub.first_header_nbytes = s->first.size();
        s->pub.nblocks = nblocks;
        s->pub.blocksizes = s->blocksizes.data();
        s->pub.data = s->data.data();
        s->pub.data_nbytes = s->data.size();
        s->pub.coded_serially = serial ? 1 : 0;
        *out = &s.release()->pub;
        return LEON_OK;
    } catch (const std::bad_alloc &) {
        leon_internal_set_error("leon_headers_encode: out of host memory");
        return LEON_E_NOMEM;
    }
}

// Inverse of leon_headers_encode: the names of every block, one per line.  Blocks are decoded in parallel with a fresh
// object each; when a block consulted state an earlier block would have left (see the top of this file) the whole
// stream is decoded again by one serial object, like the reference with -nb-cores 1.
extern "C" int leon_headers_decode(const leon_header_stream *hs, int32_t nthreads, leon_decoded **out) {
    try {
        if (!hs || !out || (hs->nblocks && (!hs->blocksizes || !hs->data))) { leon_internal_set_error("leon_headers_decode: null argument"); return LEON_E_ARG; }
        std::unique_ptr<leon_decoded> d(new leon_decoded());
        const uint64_t nb = hs->nblocks;
        std::vector<uint64_t> start(nb + 1, 0);
        for (uint64_t b = 0; b < nb; b++) start[b + 1] = start[b] + hs->blocksizes[2 * b];
        if (start[nb] > hs->data_nbytes) { leon_internal_set_error("leon_headers_decode: blocksizes exceed data_nbytes"); return LEON_E_ARG; }
        const View first{(const char *)hs->first_header, (size_t)hs->first_header_nbytes};
        std::atomic<int> leftover(0), bad(0);
        unsigned nt = nthreads > 0 ? (unsigned)nthreads : std::max(1u, std::thread::hardware_concurrency());
        nt = (unsigned)std::min<uint64_t>(nt, std::max<uint64_t>(nb, 1));
        bool serial = nt <= 1;
        if (!serial) {
            d->blocks.assign(nb, std::string());
            std::atomic<uint64_t> next(0);
            auto work = [&]() {
                for (;;) {
                    const uint64_t b = next.fetch_add(1);
                    if (b >= nb || leftover.load() || bad.load()) return;
                    HeaderDecoder h(first);
                    h.run(hs->data + start[b], hs->blocksizes[2 * b], hs->blocksizes[2 * b + 1], d->blocks[b]);
                    if (h.leftover_used) leftover.store(1);
                    if (h.bad_stream) bad.store(1);
                }
            };
            std::vector<std::thread> pool;
            for (unsigned i = 1; i < nt; i++) pool.emplace_back(work);
            work();
            for (auto &t : pool) t.join();
            serial = leftover.load() != 0;
        }
        if (serial) {
            bad.store(0);
            d->blocks.assign(nb, std::string());
            HeaderDecoder h(first);
            for (uint64_t b = 0; b < nb && !h.bad_stream; b++)
                h.run(hs->data + start[b], hs->blocksizes[2 * b], hs->blocksizes[2 * b + 1], d->blocks[b]);
            if (h.bad_stream) bad.store(1);
        }
        if (bad.load()) { leon_internal_set_error("leon_headers_decode: corrupt header stream"); return LEON_E_ARG; }
        *out = d.release();
        return LEON_OK;
    } catch (const std::bad_alloc &) {
        leon_internal_set_error("leon_headers_decode: out of host memory");
        return LEON_E_NOMEM;
    }
}

extern "C" void leon_headers_free(leon_header_stream *p) {
    if (p) delete reinterpret_cast<Stream *>(p);         // pub is the first member
}
