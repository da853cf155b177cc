// TEST INFRASTRUCTURE: the 8-bytes-per-load container skip and span compare of nbrm_json.cuh (device code
// in the product) compiled by g++ and checked against the byte-at-a-time skip on
// random JSON-like text, at every alignment.  Prints "ok <cases>" or the first
// difference; run by tests/test_wide_scanners.py.
#define NBRM_WIDE_HOST 1
#include "../../kubeflow-rm_b200/csrc/nbrm_json.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace nbrm;

// the scalar container skip, restated from skip_value's byte loop
static uint32_t scalar_skip(const char* p, uint32_t i, uint32_t n) {
  uint32_t depth = 0, j = i;
  while (j < n) {
    char c = p[j];
    if (c == '"') {
      ++j;
      for (;;) {
        if (j >= n) return n;
        if (p[j] == '\\') {
          j += 2;
        } else if (p[j] == '"') {
          ++j;
          break;
        } else {
          ++j;
        }
      }
      if (j > n) return n;
      continue;
    }
    if (c == '{' || c == '[') {
      ++depth;
    } else if (c == '}' || c == ']') {
      if (--depth == 0) return j + 1;
    }
    ++j;
  }
  return n;
}

// skip_string and skip_key, restated from the byte loops of nbrm_json.cuh
static uint32_t scalar_string(const char* p, uint32_t i, uint32_t n) {
  uint32_t j = i + 1;
  while (j < n) {
    char c = p[j];
    if (c == '\\') j += 2;
    else if (c == '"') return j + 1;
    else ++j;
  }
  return n;
}
static uint32_t scalar_key(const char* p, uint32_t i, uint32_t n, uint32_t& h) {
  uint32_t j = i + 1;
  h = 2166136261u;
  while (j < n) {
    char c = p[j];
    if (c == '"') return j + 1;
    h = (h ^ (uint32_t)(unsigned char)c) * 16777619u;
    if (c == '\\' && j + 1 < n) {
      h = (h ^ (uint32_t)(unsigned char)p[j + 1]) * 16777619u;
      j += 2;
    } else {
      ++j;
    }
  }
  return n;
}

static uint64_t rng_state = 0x9e3779b97f4a7c15ull;
static uint32_t rnd() {
  rng_state ^= rng_state << 13;
  rng_state ^= rng_state >> 7;
  rng_state ^= rng_state << 17;
  return (uint32_t)(rng_state >> 32);
}

static void gen_value(std::string& s, int depth);

static void gen_string(std::string& s) {
  s += '"';
  uint32_t len = rnd() % 14;
  for (uint32_t k = 0; k < len; ++k) {
    uint32_t r = rnd() % 40;
    if (r == 0) s += "\\\"";
    else if (r == 1) s += "\\\\";
    else if (r == 2) s += '{';
    else if (r == 3) s += ']';
    else if (r == 4) s += "\\n";
    else s += (char)('a' + r % 26);
  }
  s += '"';
}

static void gen_value(std::string& s, int depth) {
  uint32_t r = rnd() % 10;
  if (depth > 4) r %= 5;
  if (r < 3) {
    gen_string(s);
  } else if (r < 5) {
    s += std::to_string(rnd() % 100000);
  } else if (r < 8) {
    s += '{';
    uint32_t m = rnd() % 5;
    for (uint32_t k = 0; k < m; ++k) {
      if (k) s += (rnd() % 4 == 0) ? ", " : ",";
      gen_string(s);
      s += ':';
      gen_value(s, depth + 1);
    }
    s += '}';
  } else {
    s += '[';
    uint32_t m = rnd() % 5;
    for (uint32_t k = 0; k < m; ++k) {
      if (k) s += ',';
      gen_value(s, depth + 1);
    }
    s += ']';
  }
}

int main() {
  unsigned long cases = 0;
  std::vector<char> buf(1 << 16);
  for (int iter = 0; iter < 4000; ++iter) {
    std::string doc;
    doc += (rnd() & 1) ? '{' : '[';
    uint32_t m = 1 + rnd() % 6;
    bool obj = doc[0] == '{';
    for (uint32_t k = 0; k < m; ++k) {
      if (k) doc += ',';
      if (obj) {
        gen_string(doc);
        doc += ':';
      }
      gen_value(doc, 1);
    }
    doc += obj ? '}' : ']';
    std::string tail = ",\"x\":{\"y\":[1,2]}}";
    for (uint32_t align = 0; align < 8; ++align) {
      // a 16-byte aligned buffer, the document at every offset, garbage around it
      char* base = (char*)(((uintptr_t)buf.data() + 15) & ~(uintptr_t)15);
      memset(base, '}', 64);
      char* p = base + 16 + align;
      std::string text = doc + tail;
      memcpy(p, text.data(), text.size());
      memset(p + text.size(), '"', 16);
      // whole text, and truncations (unterminated containers / strings / escapes)
      uint32_t lens[4] = {(uint32_t)text.size(), (uint32_t)doc.size(),
                          (uint32_t)(rnd() % (doc.size() + 1)), (uint32_t)(doc.size() - 1)};
      for (uint32_t n : lens) {
        if (n == 0) continue;
        const uint32_t want = scalar_skip(p, 0, n);
        const uint32_t got = wide::skip_container(p, 0, n);
        ++cases;
        if (want != got) {
          printf("skip_container differs: want %u got %u n %u align %u doc %.*s\n", want, got, n,
                 align, (int)n, p);
          return 1;
        }
      }
      // inner containers, started mid-document
      for (uint32_t i = 1; i < doc.size(); ++i) {
        if (doc[i] != '{' && doc[i] != '[') continue;
        // only real container starts: the scalar walk from 0 is outside a string here
        bool in_str = false;
        for (uint32_t k = 0; k < i; ++k) {
          if (doc[k] == '\\' && in_str) ++k;
          else if (doc[k] == '"') in_str = !in_str;
        }
        if (in_str) continue;
        const uint32_t n = (uint32_t)doc.size();
        const uint32_t want = scalar_skip(p, i, n);
        const uint32_t got = wide::skip_container(p, i, n);
        ++cases;
        if (want != got) {
          printf("inner skip differs at %u: want %u got %u align %u doc %s\n", i, want, got, align,
                 doc.c_str());
          return 1;
        }
      }
      // the string scanners from every quote of the text (whatever it opens), whole and truncated
      for (uint32_t i = 0; i < text.size(); ++i) {
        if (p[i] != '"') continue;
        const uint32_t ns[3] = {(uint32_t)text.size(), i + 1 + rnd() % 12, i + 1};
        for (uint32_t n : ns) {
          if (n > text.size() || n <= i) continue;
          const uint32_t want = scalar_string(p, i, n), got = wide::skip_string(p, i, n);
          uint32_t hw = 0, hg = 0;
          const uint32_t wantk = scalar_key(p, i, n, hw), gotk = wide::skip_key(p, i, n, hg);
          cases += 2;
          if (want != got || wantk != gotk || (wantk < n && hw != hg) || (wantk == n && gotk == n && hw != hg)) {
            printf("string scan differs at %u n %u align %u: string %u/%u key %u/%u hash %08x/%08x text %.*s\n", i, n,
                   align, want, got, wantk, gotk, hw, hg, (int)n, p);
            return 1;
          }
        }
      }
      // load8: up to eight bytes from any alignment
      for (uint32_t o = 0; o + 8 <= text.size() && o < 24; ++o)
        for (uint32_t nb = 1; nb <= 8; ++nb) {
          uint64_t want = 0;
          memcpy(&want, p + o, nb);
          uint64_t got = wide::load8(p + o, nb);
          if (nb < 8) got &= (1ull << (8 * nb)) - 1;
          ++cases;
          if (want != got) {
            printf("load8 differs at %u nb %u align %u\n", o, nb, align);
            return 1;
          }
        }
      // bytes_equal against the obvious loop, both operands at every alignment
      for (uint32_t balign = 0; balign < 8; ++balign) {
        char* q = base + 32768 + balign;
        memcpy(q, doc.data(), doc.size());
        const uint32_t n = (uint32_t)doc.size();
        q[n] = (char)(p[n] ^ 1);  // a difference just past the span must not matter
        ++cases;
        if (!wide::bytes_equal(p, q, n)) {
          printf("bytes_equal false on equal spans n %u align %u/%u\n", n, align, balign);
          return 1;
        }
        const uint32_t flip = rnd() % n;
        q[flip] ^= 1;
        ++cases;
        if (wide::bytes_equal(p, q, n)) {
          printf("bytes_equal true with byte %u flipped n %u align %u/%u\n", flip, n, align, balign);
          return 1;
        }
      }
    }
  }
  printf("ok %lu\n", cases);
  return 0;
}
