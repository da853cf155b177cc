// kfam.cc — the RoleBinding name kfam gives a contributor binding.
// Reference: components/access-management/kfam/bindings.go:61-78 getBindingName, pinned by
// bindings_test.go:25-41 (tests/test_host_store.py carries the table).
#include <stdlib.h>
#include <string.h>

#include <string>

#include "../../include/nbrm_host.h"

namespace {
bool alnum(unsigned char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9'); }
// regexp.MustCompile("[^a-zA-Z0-9]+").ReplaceAllString(s, "-"); bytes of a multi-byte rune are
// all outside the class and runs collapse, so bytes and runes give the same result
std::string squeeze(const std::string& s) {
  std::string o;
  bool in_run = false;
  for (unsigned char c : s) {
    if (alnum(c)) {
      o += (char)c;
      in_run = false;
    } else if (!in_run) {
      o += '-';
      in_run = true;
    }
  }
  return o;
}
}  // namespace

extern "C" char* nbrm_kfam_binding_name(const char* user_kind, const char* user_name, const char* roleref_kind,
                                        const char* roleref_name) {
  std::string raw = std::string(user_kind) + "-" + squeeze(user_name) + "-" + roleref_kind + "-" + roleref_name;
  for (char& c : raw)
    if (c >= 'A' && c <= 'Z') c = (char)(c - 'A' + 'a');  // strings.ToLower (ASCII: the rest is squeezed away)
  const std::string out = squeeze(raw);
  char* p = (char*)malloc(out.size() + 1);
  memcpy(p, out.c_str(), out.size() + 1);
  return p;
}
