// b200_framing.cpp — see b200_framing.hpp.
#include "b200_framing.hpp"

#include <string.h>

namespace b200 {
namespace framing {

namespace {
const char kStreamId[10] = {(char)0xff, 0x06, 0x00, 0x00, 's', 'N', 'a', 'P', 'p', 'Y'};  // 4.1

void put_header(std::string* s, uint8_t type, size_t len) {
    char h[4] = {(char)type, (char)(len & 0xff), (char)((len >> 8) & 0xff), (char)((len >> 16) & 0xff)};
    s->append(h, 4);
}
void put_le32(std::string* s, uint32_t v) {
    char b[4] = {(char)(v & 0xff), (char)((v >> 8) & 0xff), (char)((v >> 16) & 0xff), (char)(v >> 24)};
    s->append(b, 4);
}

// flatten a block list (framed streams are parsed sequentially on the host; only chunk data goes to the GPU)
std::string flatten(const BlockList& b) {
    std::string s;
    for (const b2c_iovec& v : b) s.append(static_cast<const char*>(v.iov_base), v.iov_len);
    return s;
}
}  // namespace

void EncodeBatch(const std::vector<BlockList>& payloads, std::vector<std::string>* framed) {
    const size_t n = payloads.size();
    framed->assign(n, std::string());
    // pieces of every payload; a piece may span block boundaries, so flatten once per payload
    std::vector<std::string> flat(n);
    std::vector<BlockList> pieces;
    std::vector<size_t> first(n + 1, 0);
    for (size_t i = 0; i < n; i++) {
        flat[i] = flatten(payloads[i]);
        first[i] = pieces.size();
        for (size_t p = 0; p < flat[i].size(); p += kMaxChunkData)
            pieces.push_back({b2c_iovec{flat[i].data() + p, std::min(kMaxChunkData, flat[i].size() - p)}});
    }
    first[n] = pieces.size();
    std::vector<std::string> z;
    std::vector<uint32_t> crc;
    policy::SnappyCompressBatch(pieces, &z);
    policy::Crc32cBatch(pieces, nullptr, &crc);
    for (size_t i = 0; i < n; i++) {
        std::string& out = (*framed)[i];
        out.assign(kStreamId, sizeof(kStreamId));
        for (size_t k = first[i]; k < first[i + 1]; k++) {
            const size_t raw = pieces[k][0].iov_len;
            const bool compressed = z[k].size() < raw;
            const size_t data = compressed ? z[k].size() : raw;
            put_header(&out, compressed ? 0x00 : 0x01, data + 4);
            put_le32(&out, crc32c::Mask(crc[k]));
            if (compressed) out.append(z[k]);
            else out.append(static_cast<const char*>(pieces[k][0].iov_base), raw);
        }
    }
}

void DecodeBatch(const std::vector<BlockList>& framed, std::vector<std::string>* payloads, std::vector<bool>* ok) {
    const size_t n = framed.size();
    payloads->assign(n, std::string());
    ok->assign(n, true);
    struct Chunk {
        size_t stream;
        bool compressed;
        const char* data;
        size_t len;
        uint32_t masked_crc;
    };
    std::vector<std::string> flat(n);
    std::vector<Chunk> chunks;
    for (size_t i = 0; i < n; i++) {
        flat[i] = flatten(framed[i]);
        const std::string& s = flat[i];
        // "The first chunk must always be the stream identifier chunk"
        if (s.size() < sizeof(kStreamId) || memcmp(s.data(), kStreamId, sizeof(kStreamId)) != 0) {
            (*ok)[i] = false;
            continue;
        }
        size_t p = sizeof(kStreamId);
        while (p < s.size()) {
            if (s.size() - p < 4) { (*ok)[i] = false; break; }
            const uint8_t type = (uint8_t)s[p];
            const size_t len = (uint8_t)s[p + 1] | ((size_t)(uint8_t)s[p + 2] << 8) | ((size_t)(uint8_t)s[p + 3] << 16);
            p += 4;
            if (s.size() - p < len) { (*ok)[i] = false; break; }  // chunk cut short
            if (type == 0x00 || type == 0x01) {
                if (len < 4 || (type == 0x01 && len - 4 > kMaxChunkData)) { (*ok)[i] = false; break; }
                const uint8_t* c = reinterpret_cast<const uint8_t*>(s.data() + p);
                const uint32_t masked = c[0] | (c[1] << 8) | (c[2] << 16) | ((uint32_t)c[3] << 24);
                chunks.push_back({i, type == 0x00, s.data() + p + 4, len - 4, masked});
            } else if (type == 0xff) {  // repeated stream identifier: must have the right length and contents
                if (len != 6 || memcmp(s.data() + p, kStreamId + 4, 6) != 0) { (*ok)[i] = false; break; }
            } else if (type >= 0x02 && type <= 0x7f) {  // 4.5 reserved unskippable
                (*ok)[i] = false;
                break;
            }  // 0xfe padding and 0x80-0xfd reserved skippable: ignore
            p += len;
        }
    }
    // decompress every compressed chunk of every (still valid) stream in one batch
    std::vector<BlockList> zin;
    std::vector<size_t> zidx;
    for (size_t k = 0; k < chunks.size(); k++) {
        if (!(*ok)[chunks[k].stream] || !chunks[k].compressed) continue;
        uint32_t ulen = 0;
        if (!b2c_snappy_read_preamble(chunks[k].data, chunks[k].len, &ulen) || ulen > kMaxChunkData) {
            (*ok)[chunks[k].stream] = false;  // "the uncompressed data in a chunk must be no longer than 65536 bytes"
            continue;
        }
        zin.push_back({b2c_iovec{chunks[k].data, chunks[k].len}});
        zidx.push_back(k);
    }
    std::vector<std::string> raw(chunks.size());
    {
        std::vector<std::string> out;
        std::vector<bool> good;
        policy::SnappyDecompressBatch(zin, &out, &good);
        for (size_t j = 0; j < zidx.size(); j++) {
            if (!good[j]) (*ok)[chunks[zidx[j]].stream] = false;
            else raw[zidx[j]].swap(out[j]);
        }
    }
    // checksum every chunk's uncompressed bytes in one batch
    std::vector<BlockList> cin;
    std::vector<size_t> cidx;
    for (size_t k = 0; k < chunks.size(); k++) {
        if (!(*ok)[chunks[k].stream]) continue;
        if (chunks[k].compressed) cin.push_back({b2c_iovec{raw[k].data(), raw[k].size()}});
        else cin.push_back({b2c_iovec{chunks[k].data, chunks[k].len}});
        cidx.push_back(k);
    }
    std::vector<uint32_t> crc;
    policy::Crc32cBatch(cin, nullptr, &crc);
    for (size_t j = 0; j < cidx.size(); j++)
        if (crc32c::Mask(crc[j]) != chunks[cidx[j]].masked_crc) (*ok)[chunks[cidx[j]].stream] = false;
    for (size_t k = 0; k < chunks.size(); k++) {
        const size_t i = chunks[k].stream;
        if (!(*ok)[i]) continue;
        if (chunks[k].compressed) (*payloads)[i].append(raw[k]);
        else (*payloads)[i].append(chunks[k].data, chunks[k].len);
    }
    for (size_t i = 0; i < n; i++)
        if (!(*ok)[i]) (*payloads)[i].clear();
}

}  // namespace framing
}  // namespace b200
