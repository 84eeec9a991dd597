// TEST DOUBLE of butil/iobuf.h: just enough of butil::IOBuf for brpc_gpu_snappy.cpp to compile and run outside a
// bRPC tree.  Same member names, argument meaning and return values as the reference for the calls the shim makes
// (src/butil/iobuf.h:74-101, 360-363; src/butil/iobuf.cpp:1166-1188, 1216-1240, 1390-1413, 1504-1510): an IOBuf is
// a list of references to blocks; append() copies into 8160-byte payload blocks (an 8192-byte Block minus its
// header, iobuf.h:67); append_user_data_with_meta() adopts the caller's buffer and runs its deleter when the last
// reference goes.  Not a re-implementation of IOBuf.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include <memory>
#include <string>
#include <vector>

namespace butil {

class StringPiece {
public:
    StringPiece(const char* p, size_t n) : p_(p), n_(n) {}
    const char* data() const { return p_; }
    size_t size() const { return n_; }

private:
    const char* p_;
    size_t n_;
};

class IOBuf {
public:
    static const size_t DEFAULT_PAYLOAD = 8160;
    struct Block {
        char* data;
        size_t cap, size;
        void (*deleter)(void*);
        uint64_t meta;
        ~Block() {
            if (deleter) deleter(data);
            else delete[] data;
        }
    };
    struct Ref {
        std::shared_ptr<Block> b;
        size_t off, len;
    };

    size_t size() const {
        size_t n = 0;
        for (const Ref& r : refs_) n += r.len;
        return n;
    }
    bool empty() const { return size() == 0; }
    void clear() { refs_.clear(); }
    size_t backing_block_num() const { return refs_.size(); }
    StringPiece backing_block(size_t i) const { return StringPiece(refs_[i].b->data + refs_[i].off, refs_[i].len); }
    int append(const void* data, size_t n) {
        const char* p = static_cast<const char*>(data);
        while (n) {
            if (refs_.empty() || refs_.back().b->deleter || refs_.back().b->size == refs_.back().b->cap ||
                refs_.back().off + refs_.back().len != refs_.back().b->size) {
                std::shared_ptr<Block> b(new Block{new char[DEFAULT_PAYLOAD], DEFAULT_PAYLOAD, 0, nullptr, 0});
                refs_.push_back(Ref{b, 0, 0});
            }
            Ref& r = refs_.back();
            const size_t c = n < r.b->cap - r.b->size ? n : r.b->cap - r.b->size;
            memcpy(r.b->data + r.b->size, p, c);
            r.b->size += c;
            r.len += c;
            p += c;
            n -= c;
        }
        return 0;
    }
    int append(const std::string& s) { return append(s.data(), s.size()); }
    int append_user_data_with_meta(void* data, size_t size, void (*deleter)(void*), uint64_t meta) {
        if (size > 0xFFFFFFFFULL - 100) return -1;
        if (!deleter) deleter = ::free;
        if (!size) {
            deleter(data);
            return 0;
        }
        std::shared_ptr<Block> b(new Block{static_cast<char*>(data), size, size, deleter, meta});
        refs_.push_back(Ref{b, 0, size});
        return 0;
    }
    int append_user_data(void* data, size_t size, void (*deleter)(void*)) {
        return append_user_data_with_meta(data, size, deleter, 0);
    }
    uint64_t get_first_data_meta() const { return refs_.empty() ? 0 : refs_[0].b->meta; }
    size_t copy_to(void* buf, size_t n = (size_t)-1, size_t pos = 0) const {
        char* o = static_cast<char*>(buf);
        size_t done = 0;
        for (const Ref& r : refs_) {
            if (pos >= r.len) {
                pos -= r.len;
                continue;
            }
            const size_t c = n - done < r.len - pos ? n - done : r.len - pos;
            memcpy(o + done, r.b->data + r.off + pos, c);
            done += c;
            pos = 0;
            if (done == n) break;
        }
        return done;
    }
    std::string to_string() const {
        std::string s(size(), '\0');
        if (!s.empty()) copy_to(&s[0], s.size());
        return s;
    }

private:
    std::vector<Ref> refs_;
};

class IOBufAsZeroCopyOutputStream {
public:
    explicit IOBufAsZeroCopyOutputStream(IOBuf* buf) : buf_(buf) {}
    IOBuf* buf() const { return buf_; }

private:
    IOBuf* buf_;
};

}  // namespace butil
