// b200_framing.hpp — the snappy framing format (".sz", application/x-snappy-framed) on top of the
// GPU batch codec.  The reference vendors the specification but no implementation
// (src/butil/third_party/snappy/framing_format.txt); this follows that text section by section.
// Every <= 64 KiB chunk (4.2/4.3) is an independent snappy stream with a masked CRC-32C of its
// uncompressed bytes (section 3) — exactly the unit the batch kernels parallelise over, so all
// chunks of all payloads of a call go to the GPU as one batch.
#pragma once

#include <string>
#include <vector>

#include "b200_codec.hpp"

namespace b200 {
namespace framing {

constexpr size_t kMaxChunkData = 65536;  // framing_format.txt 4.2 / 4.3

// Encode each payload as one framed stream: stream identifier (4.1) followed by one chunk per
// 64 KiB piece — compressed (type 0x00) when that is smaller than the piece, else uncompressed
// (type 0x01).
void EncodeBatch(const std::vector<BlockList>& payloads, std::vector<std::string>* framed);

// Decode framed streams.  ok[i] is false for a stream that violates the format: missing or
// malformed stream identifier, reserved unskippable chunk (4.5), chunk cut short, chunk data that
// is not a valid snappy stream or decodes to more than 64 KiB, checksum mismatch.  Padding (4.4),
// reserved skippable chunks (4.6) and repeated stream identifiers are skipped.
void DecodeBatch(const std::vector<BlockList>& framed, std::vector<std::string>* payloads, std::vector<bool>* ok);

inline void Encode(const BlockList& payload, std::string* framed) {
    std::vector<std::string> out;
    EncodeBatch({payload}, &out);
    framed->append(out[0]);
}
inline bool Decode(const BlockList& framed, std::string* payload) {
    std::vector<std::string> out;
    std::vector<bool> ok;
    DecodeBatch({framed}, &out, &ok);
    if (ok[0]) payload->append(out[0]);
    return ok[0];
}

}  // namespace framing
}  // namespace b200
