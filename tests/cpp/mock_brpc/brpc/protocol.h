// TEST DOUBLE of brpc/protocol.h: ParsePbFromIOBuf (src/brpc/protocol.cpp).
#pragma once
#include <google/protobuf/message.h>

#include "butil/iobuf.h"

namespace brpc {
inline bool ParsePbFromIOBuf(google::protobuf::Message* msg, const butil::IOBuf& buf) { return msg->ParseFromString(buf.to_string()); }
}  // namespace brpc
