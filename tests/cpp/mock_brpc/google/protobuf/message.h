// TEST DOUBLE of google/protobuf/message.h: a Message that serialises to / parses from a byte string, with the
// one stream call brpc's snappy handlers make (src/brpc/policy/snappy_compress.cpp:28-38).
#pragma once
#include <string>

#include "butil/iobuf.h"

namespace google {
namespace protobuf {

class Message {
public:
    virtual ~Message() {}
    virtual bool SerializeToString(std::string* out) const = 0;
    virtual bool ParseFromString(const std::string& in) = 0;
    bool SerializeToZeroCopyStream(butil::IOBufAsZeroCopyOutputStream* s) const {
        std::string bytes;
        if (!SerializeToString(&bytes)) return false;
        s->buf()->append(bytes);
        return true;
    }
};

}  // namespace protobuf
}  // namespace google
