// Internal: decoded text of a stream, one string per block (behind the opaque leon_decoded of include/leon_b200.h).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

struct leon_decoded {
    std::vector<std::string> blocks;
};
