// dbow3_stream.cu -- host side only: the payload of a DBoW3 binary vocabulary stream (what a .dbow3 file holds).
//
// Vocabulary::toStream (thirdParty/DBow3/src/Vocabulary.cpp:1180-1256) writes
//   u64 signature | u8 compressed | u32 node count | payload
// and, when `compressed` (its default, so what the reference's orbvoc.dbow3 is), cuts the payload into 10 000-byte
// chunks (:1228-1249), each passed through qlz_compress of the vendored QuickLZ 1.5.0 at its shipped settings
// (quicklz.h:25/31: compression level 1, streaming buffer 0), preceded by a u32 chunk count.  Vocabulary::fromStream
// (:1350-1367) inflates chunk after chunk.  This file restates the level-1 block format from the published algorithm:
//
//   block  = header, body          header = flags, compressed size, decompressed size  (u8 sizes, or u32 when flags & 2)
//   flags  : bit 0 body is compressed (else stored verbatim), bit 1 wide header, bits 2-3 level, bit 6 always set
//   body   = groups of a 32-bit control word (bit 31 = sentinel) and up to 31 items, one per control bit from bit 0:
//              0  literal byte
//              1  match: u16 = (hash << 4) | n;  n != 0: length n + 2;  n == 0: a third byte holds the length
//            the match source is NOT an offset: it is the most recent output position whose three bytes hash to `hash`,
//            in a 4096-entry table the decoder maintains by the rule the encoder used (see hash_through below).
//   the last 4 bytes of a block are always literals, and once fewer than 11 bytes remain everything is literal.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "vocabulary_stream.h"

namespace {

inline uint32_t le(const uint8_t* p, int n) {
  uint32_t v = 0;
  for (int i = 0; i < n; i++) v |= (uint32_t)p[i] << (8 * i);
  return v;
}

// one QuickLZ level-1 block at in[0, avail) -> appended to out; *used = bytes of input the block occupies
bool inflate_block(const uint8_t* in, size_t avail, std::vector<uint8_t>& out, size_t* used, std::string* why) {
  auto fail = [&](const char* msg) {
    *why = msg;
    return false;
  };
  if (avail < 3) return fail("truncated QuickLZ block header");
  const uint8_t flags = in[0];
  const int wide = (flags & 2) ? 4 : 1;
  const size_t header = 1 + 2 * (size_t)wide;
  if (avail < header) return fail("truncated QuickLZ block header");
  const size_t csize = le(in + 1, wide), dsize = le(in + 1 + wide, wide);
  if (csize < header || csize > avail) return fail("QuickLZ block longer than the file");
  *used = csize;
  const size_t base = out.size();
  if (!(flags & 1)) {  // stored
    if (csize - header != dsize) return fail("stored QuickLZ block with inconsistent sizes");
    out.insert(out.end(), in + header, in + csize);
    return true;
  }
  if (((flags >> 2) & 3) != 1) return fail("QuickLZ block not written at compression level 1 (DBoW3's setting)");
  out.resize(base + dsize);
  uint8_t* dst = out.data() + base;
  int32_t table[4096];
  for (int i = 0; i < 4096; i++) table[i] = -1;
  long hashed = -1;  // every position <= hashed that the encoder entered into its table is in ours
  auto hash_through = [&](long last) {
    for (; hashed < last;) {
      ++hashed;
      const uint32_t v = le(dst + hashed, 3);
      table[((v >> 12) ^ v) & 0xfff] = (int32_t)hashed;
    }
  };
  const long n = (long)dsize, literal_zone = n - 11;  // at or beyond it the encoder emits literals only
  size_t sp = header;
  long pos = 0;
  uint32_t ctrl = 1;
  for (;;) {
    if (ctrl == 1) {
      if (sp + 4 > csize) return fail("QuickLZ control word past the block");
      ctrl = le(in + sp, 4);
      sp += 4;
    }
    if (ctrl & 1) {
      ctrl >>= 1;
      if (sp + 2 > csize) return fail("QuickLZ match past the block");
      const uint32_t f = le(in + sp, 2);
      long len = f & 0xf;
      if (len) {
        len += 2;
        sp += 2;
      } else {
        if (sp + 3 > csize) return fail("QuickLZ match past the block");
        len = in[sp + 2];
        sp += 3;
      }
      const long from = table[(f >> 4) & 0xfff];
      if (from < 0 || from >= pos || len < 3 || pos + len > n - 4) return fail("corrupt QuickLZ match");
      for (long i = 0; i < len; i++) dst[pos + i] = dst[from + i];  // forward, byte by byte: overlap replicates
      hash_through(pos);  // the match start itself; the encoder does not enter the positions a match covers
      pos += len;
      hashed = pos - 1;
    } else if (pos < literal_zone) {
      int run = 0;  // literals up to the next match bit, at most 4 at a time
      while (run < 4 && !((ctrl >> run) & 1)) run++;
      if (sp + run > csize || pos + run > n) return fail("QuickLZ literals past the block");
      std::memcpy(dst + pos, in + sp, (size_t)run);
      ctrl >>= run;
      pos += run;
      sp += run;
      hash_through(pos - 3);
    } else {
      while (pos < n) {
        if (ctrl == 1) {
          sp += 4;
          ctrl = 1u << 31;
        }
        if (sp >= csize) return fail("QuickLZ tail past the block");
        dst[pos++] = in[sp++];
        ctrl >>= 1;
      }
      return true;
    }
  }
}

}  // namespace

bool dbow3_inflate_chunks(const uint8_t* in, size_t avail, std::vector<uint8_t>& payload, std::string* why) {
  payload.clear();
  if (avail < 4) {
    *why = "truncated chunk count";
    return false;
  }
  const uint32_t chunks = le(in, 4);  // Vocabulary.cpp:1357-1358
  size_t at = 4;
  for (uint32_t c = 0; c < chunks; c++) {
    size_t used = 0;
    if (!inflate_block(in + at, avail - at, payload, &used, why)) return false;
    at += used;
  }
  return true;
}
