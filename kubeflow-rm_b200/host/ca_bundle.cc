// ca_bundle.cc — the pure part of the trusted-CA bundle path of the ODH reconciler.
//
// Reference: components/odh-notebook-controller/controllers/notebook_controller.go —
// CreateNotebookCertConfigMap :254-357 (which certificate files make it into
// workbench-trusted-ca-bundle, and the joined bundle), IsConfigMapDeleted :361-387,
// UnsetNotebookCertConfig :390-455 (with its delete-while-ranging loops :411-415, :418-422,
// reproduced on Go's slice semantics).  The ConfigMap Gets / Create / Update / Patch are I/O
// and stay with the caller.  Go's crypto/x509 is not restated: the first PEM block of a file
// must be a CERTIFICATE whose DER has the outer shape of one; a caller with a real X.509
// parser passes its verdicts in x509_ok.  Checked against oracle/cabundle.py, which is pinned
// by the reference's own two-certificate fixtures (tests/test_oracle_cabundle.py).
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/nbrm_host.h"
#include "../csrc/nbrm_json.cuh"

namespace {
using namespace nbrm;

bool go_space(unsigned char c) {  // unicode.IsSpace for Latin-1, as strings.TrimSpace uses it
  return c == ' ' || (c >= 9 && c <= 13) || c == 0x85 || c == 0xA0;
}
std::string trim_space(const std::string& s) {
  size_t a = 0, b = s.size();
  while (a < b && go_space((unsigned char)s[a])) ++a;
  while (b > a && go_space((unsigned char)s[b - 1])) --b;
  return s.substr(a, b - a);
}

int b64val(char c) {
  if (c >= 'A' && c <= 'Z') return c - 'A';
  if (c >= 'a' && c <= 'z') return c - 'a' + 26;
  if (c >= '0' && c <= '9') return c - '0' + 52;
  if (c == '+') return 62;
  if (c == '/') return 63;
  return -1;
}

// pem.Decode: the first block of `data` from position *pos on.  Returns false when there is none.
bool pem_decode(const std::string& data, size_t* pos, std::string* type, std::string* der) {
  size_t from = *pos;
  for (;;) {
    size_t b = data.find("-----BEGIN ", from);
    if (b == std::string::npos) return false;
    if (b != 0 && data[b - 1] != '\n') {
      from = b + 1;
      continue;
    }
    size_t te = data.find("-----", b + 11);
    size_t eol = data.find('\n', b);
    if (te == std::string::npos || eol == std::string::npos || te > eol) return false;
    *type = data.substr(b + 11, te - (b + 11));
    const std::string endline = "\n-----END " + *type + "-----";
    size_t e = data.find(endline, eol);
    if (e == std::string::npos) return false;
    std::string body = data.substr(eol + 1, e - eol);
    // headers: "Key: value" lines up to a blank line
    size_t p = 0;
    bool had_header = false;
    for (;;) {
      size_t nl = body.find('\n', p);
      std::string line = body.substr(p, nl == std::string::npos ? nl : nl - p);
      if (line.find(':') == std::string::npos) break;
      had_header = true;
      if (nl == std::string::npos) {
        p = body.size();
        break;
      }
      p = nl + 1;
    }
    if (had_header) {
      size_t nl = body.find('\n', p);
      if (trim_space(body.substr(p, nl == std::string::npos ? nl : nl - p)).empty() && nl != std::string::npos) p = nl + 1;
    }
    der->clear();
    uint32_t acc = 0;
    int bits = 0, pad = 0;
    for (size_t i = p; i < body.size(); ++i) {
      const char c = body[i];
      if (c == '\n' || c == '\r' || c == ' ' || c == '\t') continue;
      if (c == '=') {
        ++pad;
        continue;
      }
      const int v = b64val(c);
      if (v < 0 || pad) return false;
      acc = (acc << 6) | (uint32_t)v;
      bits += 6;
      if (bits >= 8) {
        bits -= 8;
        der->push_back((char)((acc >> bits) & 0xff));
      }
    }
    size_t after = e + endline.size();
    if (after < data.size() && data[after] == '\r') ++after;
    if (after < data.size() && data[after] == '\n') ++after;
    *pos = after;
    return true;
  }
}

bool der_tlv(const std::string& b, size_t i, int* tag, size_t* cs, size_t* ce) {
  if (i + 2 > b.size()) return false;
  *tag = (unsigned char)b[i];
  size_t ln = (unsigned char)b[i + 1];
  i += 2;
  if (ln & 0x80) {
    const size_t k = ln & 0x7f;
    if (k == 0 || k > 4 || i + k > b.size()) return false;
    ln = 0;
    for (size_t q = 0; q < k; ++q) ln = (ln << 8) | (unsigned char)b[i + q];
    i += k;
  }
  if (i + ln > b.size()) return false;
  *cs = i;
  *ce = i + ln;
  return true;
}

// Certificate ::= SEQUENCE { tbsCertificate SEQUENCE, signatureAlgorithm SEQUENCE, signatureValue BIT STRING }
bool der_is_certificate(const std::string& der) {
  int tag;
  size_t cs, ce;
  if (!der_tlv(der, 0, &tag, &cs, &ce) || tag != 0x30 || ce != der.size()) return false;
  size_t i = cs;
  const int want[3] = {0x30, 0x30, 0x03};
  for (int k = 0; k < 3; ++k) {
    size_t s2, e2;
    if (!der_tlv(der, i, &tag, &s2, &e2) || tag != want[k] || e2 > ce) return false;
    i = e2;
  }
  return i == ce;
}

char* dup_out(const std::string& s, uint64_t* len) {
  char* o = (char*)malloc(s.size() + 1);
  memcpy(o, s.data(), s.size());
  o[s.size()] = 0;
  if (len) *len = s.size();
  return o;
}

std::string text(const Doc& d, Span v) { return std::string(d.p + v.off, v.len); }

// `for index, x := range s { if match(x) { s = append(s[:index], s[index+1:]...) } }`: the range
// keeps the original length and reads the backing array the appends shift.  false = Go panics.
template <class Match>
bool delete_while_ranging(std::vector<std::string>& items, Match match) {
  std::vector<std::string> backing = items;
  size_t length = backing.size();
  const size_t n = items.size();
  for (size_t index = 0; index < n; ++index) {
    if (!match(backing[index])) continue;
    if (index + 1 > length) return false;  // s[index+1:] beyond len(s)
    for (size_t k = index; k + 1 < length; ++k) backing[k] = backing[k + 1];
    --length;
  }
  backing.resize(length);
  items.swap(backing);
  return true;
}

std::string name_of(const std::string& elem) {
  Doc d;
  d.p = elem.data();
  d.n = (uint32_t)elem.size();
  const Span n = find(d, root(d), "name");
  return is_string(d, n) ? text(d, inner(n)) : std::string();
}
}  // namespace

extern "C" {

int nbrm_ca_bundle(const char* ca_bundle_crt, const char* odh_ca_bundle_crt, int odh_configmap_found,
                   const char* ca_crt, int kube_configmap_found, const int8_t* x509_ok, char** out,
                   uint64_t* len) {
  *out = nullptr;
  if (len) *len = 0;
  std::vector<std::string> pool;
  struct File {
    const char* data;
    bool is_ca_bundle;
    int slot;
  };
  const File odh[2] = {{ca_bundle_crt, true, 0}, {odh_ca_bundle_crt, false, 1}};
  const File kube[1] = {{ca_crt, false, 2}};
  for (int cm = 0; cm < 2; ++cm) {
    const bool found = cm == 0 ? odh_configmap_found != 0 : kube_configmap_found != 0;
    if (!found) {
      if (cm == 0) return 0;  // :277-279
      continue;               // :280-281
    }
    const File* files = cm == 0 ? odh : kube;
    for (int f = 0; f < (cm == 0 ? 2 : 1); ++f) {
      const bool ok = files[f].data != nullptr;
      const std::string data = trim_space(ok ? files[f].data : "");  // :289
      if (!ok || (files[f].is_ca_bundle && data.empty())) return 0;  // :293-295
      if (data.empty()) continue;                                    // :296-298
      size_t pos = 0;
      std::string type, der;
      if (pem_decode(data, &pos, &type, &der) && type == "CERTIFICATE") {  // :301-302
        const bool good = x509_ok && x509_ok[files[f].slot] >= 0 ? x509_ok[files[f].slot] != 0 : der_is_certificate(der);
        if (good) pool.push_back(data);  // :310: the whole file, not only its first block
      }
    }
  }
  if (pool.empty()) return 0;
  std::string joined;
  for (size_t i = 0; i < pool.size(); ++i) joined += (i ? "\n" : "") + pool[i];  // :325
  *out = dup_out(joined, len);
  return 1;
}

uint64_t nbrm_pem_count_certificates(const char* bundle, uint64_t len) {
  const std::string data(bundle, len);
  size_t pos = 0;
  uint64_t n = 0;
  std::string type, der;
  while (pos < data.size() && pem_decode(data, &pos, &type, &der))
    if (type == "CERTIFICATE" && der_is_certificate(der)) ++n;
  return n;
}

// UnsetNotebookCertConfig :390-455 on the Notebook's JSON.  Returns 0 nothing to change, 1 *out
// is the Notebook to patch to, -1 the reference panics (slice bounds out of range in the
// delete-while-ranging loop), -2 not a Notebook object.
int nbrm_unset_notebook_cert_config(const char* notebook, uint64_t nlen, char** out, uint64_t* len) {
  *out = nullptr;
  if (len) *len = 0;
  const std::string nb(notebook, nlen);
  Doc d;
  d.p = nb.data();
  d.n = (uint32_t)nb.size();
  const Span rt = root(d);
  if (!is_object(d, rt)) return -2;
  const Span nm = find(d, find(d, rt, "metadata"), "name");
  const std::string name = is_string(d, nm) ? text(d, inner(nm)) : std::string();
  const Span podspec = find(d, find(d, find(d, rt, "spec"), "template"), "spec");
  if (!is_object(d, podspec)) return 0;
  const Span ctrs = find(d, podspec, "containers"), vols = find(d, podspec, "volumes");
  bool changed = false;
  std::string new_ctrs, new_vols;
  {
    std::vector<std::string> elems;
    ArrIter it(d, ctrs);
    Span e;
    while (it.next(d, e)) elems.push_back(text(d, e));
    for (size_t i = 0; i < elems.size(); ++i) {
      if (name_of(elems[i]) != name) continue;  // :408
      Doc c;
      c.p = elems[i].data();
      c.n = (uint32_t)elems[i].size();
      std::string rebuilt = "{";
      bool first = true;
      ObjIter ci(c, root(c));
      Member m;
      while (ci.next(c, m)) {
        std::string val = text(c, m.val);
        const std::string key = text(c, inner(m.key));
        if ((key == "env" || key == "volumeMounts") && is_array(c, m.val)) {
          std::vector<std::string> items;
          ArrIter ai(c, m.val);
          Span a;
          while (ai.next(c, a)) items.push_back(text(c, a));
          if (key == "env") {
            const char* vars[5] = {"PIP_CERT", "REQUESTS_CA_BUNDLE", "SSL_CERT_FILE", "PIPELINES_SSL_SA_CERTS", "GIT_SSL_CAINFO"};
            for (const char* v : vars)  // :410-416
              if (!delete_while_ranging(items, [&](const std::string& x) { return name_of(x) == v; })) return -1;
          } else if (!delete_while_ranging(items, [&](const std::string& x) { return name_of(x) == "trusted-ca"; })) {  // :418-422
            return -1;
          }
          val = "[";
          for (size_t k = 0; k < items.size(); ++k) val += (k ? "," : "") + items[k];
          val += "]";
        }
        if (!first) rebuilt += ',';
        first = false;
        rebuilt += text(c, m.key) + ":" + val;
      }
      elems[i] = rebuilt + "}";
      changed = true;  // :424-430: set whenever the image container exists
      break;
    }
    new_ctrs = "[";
    for (size_t k = 0; k < elems.size(); ++k) new_ctrs += (k ? "," : "") + elems[k];
    new_ctrs += "]";
  }
  {
    std::vector<std::string> elems;
    ArrIter it(d, vols);
    Span e;
    while (it.next(d, e)) elems.push_back(text(d, e));
    for (size_t i = 0; i < elems.size(); ++i) {  // :436-442
      Doc v;
      v.p = elems[i].data();
      v.n = (uint32_t)elems[i].size();
      const Span cm = find(v, root(v), "configMap");
      const Span cn = find(v, cm, "name");
      if (is_object(v, cm) && is_string(v, cn) && text(v, inner(cn)) == "workbench-trusted-ca-bundle") {
        elems.erase(elems.begin() + i);
        changed = true;
        break;
      }
    }
    new_vols = "[";
    for (size_t k = 0; k < elems.size(); ++k) new_vols += (k ? "," : "") + elems[k];
    new_vols += "]";
  }
  if (!changed) return 0;
  // splice the two arrays back (the later span first, so the earlier offsets stay valid)
  std::string res = nb;
  struct Rep {
    Span s;
    const std::string* v;
  };
  Rep reps[2] = {{ctrs, &new_ctrs}, {vols, &new_vols}};
  if (reps[0].s.off < reps[1].s.off) std::swap(reps[0], reps[1]);
  for (auto& r : reps)
    if (is_array(d, r.s)) res.replace(r.s.off, r.s.len, *r.v);
  *out = dup_out(res, len);
  return 1;
}

}  // extern "C"
