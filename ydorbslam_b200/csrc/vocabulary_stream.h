// vocabulary_stream.h -- host helpers of the DBoW3 binary vocabulary stream (dbow3_stream.cu)
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

// `in` points just behind the 13-byte stream header of a compressed stream (at the u32 chunk count,
// Vocabulary.cpp:1357): inflates every QuickLZ level-1 chunk into `payload`.  false + *why on a malformed stream.
bool dbow3_inflate_chunks(const uint8_t* in, size_t avail, std::vector<uint8_t>& payload, std::string* why);
