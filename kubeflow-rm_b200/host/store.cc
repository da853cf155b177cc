// store.cc — see store.h.
#include "store.h"

#include <stdlib.h>

#include "../csrc/nbrm_json.cuh"

using namespace nbrm;

namespace nbrm_host {

namespace {
Doc doc_of(const std::string& s) {
  Doc d;
  d.p = s.data();
  d.n = (uint32_t)s.size();
  return d;
}
std::string unquote(const Doc& d, Span v) {
  if (!is_string(d, v)) return std::string();
  return std::string(d.p + v.off + 1, v.len - 2);  // names, kinds, versions carry no escapes
}
std::string text(const Doc& d, Span v) { return std::string(d.p + v.off, v.len); }

// json with the members of object `obj` rewritten: member `key` replaced by raw
// (dropped when raw is empty), appended when missing.
std::string rewrite_member(const std::string& json, Span obj, const std::string& key,
                           const std::string& raw) {
  Doc d = doc_of(json);
  std::string out(json, 0, obj.off);
  out += '{';
  bool first = true, done = false;
  ObjIter it(d, obj);
  Member m;
  while (it.next(d, m)) {
    const bool hit = text(d, inner(m.key)) == key;
    if (hit && (done || raw.empty())) {
      done = true;
      continue;
    }
    if (!first) out += ',';
    first = false;
    if (hit) {
      out += text(d, m.key) + ":" + raw;
      done = true;
    } else {
      out += text(d, m.all);
    }
  }
  if (!done && !raw.empty()) {
    if (!first) out += ',';
    out += "\"" + key + "\":" + raw;
  }
  out += '}';
  out.append(json, obj.end(), std::string::npos);
  return out;
}
}  // namespace

ObjMeta read_meta(const std::string& json) {
  ObjMeta r;
  Doc d = doc_of(json);
  Span rt = root(d);
  if (!is_object(d, rt)) return r;
  r.kind = unquote(d, find(d, rt, "kind"));
  Span meta = find(d, rt, "metadata");
  r.ns = unquote(d, find(d, meta, "namespace"));
  r.name = unquote(d, find(d, meta, "name"));
  r.resource_version = unquote(d, find(d, meta, "resourceVersion"));
  r.uid = unquote(d, find(d, meta, "uid"));
  r.has_deletion_timestamp = !is_nil(d, find(d, meta, "deletionTimestamp"));
  r.ok = !r.kind.empty() && !r.name.empty();
  return r;
}

std::string read_string(const std::string& json, const std::vector<std::string>& path) {
  Doc d = doc_of(json);
  Span cur = root(d);
  for (const std::string& key : path) {
    Span next = mk(0, 0);
    ObjIter it(d, cur);
    Member m;
    while (it.next(d, m))
      if (text(d, inner(m.key)) == key) next = m.val;
    cur = next;
  }
  return unquote(d, cur);
}

std::string read_annotation(const std::string& json, const std::string& key) {
  Doc d = doc_of(json);
  Span ann = find(d, find(d, root(d), "metadata"), "annotations");
  ObjIter it(d, ann);
  Member m;
  while (it.next(d, m))
    if (text(d, inner(m.key)) == key) return text(d, m.val);
  return std::string();
}

std::string with_meta_field(const std::string& json, const std::string& field, const std::string& raw) {
  Doc d = doc_of(json);
  Span rt = root(d);
  Span meta = find(d, rt, "metadata");
  if (!is_object(d, meta)) return rewrite_member(json, rt, "metadata", "{\"" + field + "\":" + raw + "}");
  return rewrite_member(json, meta, field, raw);
}

std::string with_status(const std::string& json, const std::string& raw) {
  Doc d = doc_of(json);
  return rewrite_member(json, root(d), "status", raw);
}

std::string without_annotation(const std::string& json, const std::string& key) {
  Doc d = doc_of(json);
  Span ann = find(d, find(d, root(d), "metadata"), "annotations");
  if (!is_object(d, ann)) return json;
  return rewrite_member(json, ann, key, "");
}

namespace {
std::string merge_value(const Doc& t, Span tv, const Doc& p, Span pv) {
  if (!is_object(p, pv)) return text(p, pv);  // anything but an object replaces the target
  std::string out = "{";
  bool first = true;
  auto emit = [&](const std::string& key_text, const std::string& val) {
    if (!first) out += ',';
    first = false;
    out += key_text + ":" + val;
  };
  const bool tobj = is_object(t, tv);
  if (tobj) {  // target members, patched or kept
    ObjIter it(t, tv);
    Member m;
    while (it.next(t, m)) {
      Span pm = mk(0, 0);
      ObjIter pi(p, pv);
      Member q;
      while (pi.next(p, q))
        if (text(p, inner(q.key)) == text(t, inner(m.key))) pm = q.val;
      if (!pm.ok()) emit(text(t, m.key), text(t, m.val));
      else if (is_null(p, pm)) continue;
      else emit(text(t, m.key), merge_value(t, m.val, p, pm));
    }
  }
  ObjIter pi(p, pv);  // patch members the target does not have
  Member q;
  while (pi.next(p, q)) {
    if (is_null(p, q.val)) continue;
    bool have = false;
    if (tobj) {
      ObjIter it(t, tv);
      Member m;
      while (it.next(t, m))
        if (text(p, inner(q.key)) == text(t, inner(m.key))) have = true;
    }
    if (!have) emit(text(p, q.key), merge_value(t, mk(0, 0), p, q.val));
  }
  return out + "}";
}

std::string unescape_pointer(std::string s) {
  std::string o;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '~' && i + 1 < s.size()) {
      o += s[i + 1] == '1' ? '/' : '~';
      ++i;
    } else {
      o += s[i];
    }
  }
  return o;
}

// `doc` with the value at tokens[k..] added / replaced by `val` (op 0 add, 2 replace) or removed (op 1)
bool patch_at(const Doc& d, Span v, const std::vector<std::string>& tokens, size_t k, int op,
              const std::string& val, std::string* out) {
  const std::string& tok = tokens[k];
  const bool last = k + 1 == tokens.size();
  if (is_object(d, v)) {
    *out = "{";
    bool first = true, hit = false;
    ObjIter it(d, v);
    Member m;
    while (it.next(d, m)) {
      std::string mv = text(d, m.val);
      if (text(d, inner(m.key)) == tok) {
        hit = true;
        if (last) {
          if (op == 1) continue;
          mv = val;
        } else if (!patch_at(d, m.val, tokens, k + 1, op, val, &mv)) {
          return false;
        }
      }
      if (!first) *out += ',';
      first = false;
      *out += text(d, m.key) + ":" + mv;
    }
    if (!hit) {
      if (!last || op != 0) return false;  // only "add" may name a member that is not there
      if (!first) *out += ',';
      *out += "\"" + tok + "\":" + val;
    }
    *out += "}";
    return true;
  }
  if (is_array(d, v)) {
    std::vector<std::string> elems;
    std::vector<Span> spans;
    ArrIter it(d, v);
    Span e;
    while (it.next(d, e)) {
      elems.push_back(text(d, e));
      spans.push_back(e);
    }
    if (tok == "-") {
      if (!last || op != 0) return false;
      elems.push_back(val);
    } else {
      char* endp = nullptr;
      const unsigned long idx = strtoul(tok.c_str(), &endp, 10);
      if (tok.empty() || *endp) return false;
      if (last) {
        if (op == 0 && idx <= elems.size()) elems.insert(elems.begin() + idx, val);
        else if (op == 1 && idx < elems.size()) elems.erase(elems.begin() + idx);
        else if (op == 2 && idx < elems.size()) elems[idx] = val;
        else return false;
      } else {
        if (idx >= elems.size() || !patch_at(d, spans[idx], tokens, k + 1, op, val, &elems[idx])) return false;
      }
    }
    *out = "[";
    for (size_t i = 0; i < elems.size(); ++i) *out += (i ? "," : "") + elems[i];
    *out += "]";
    return true;
  }
  return false;
}
}  // namespace

std::string apply_merge_patch(const std::string& target, const std::string& patch) {
  Doc t = doc_of(target), p = doc_of(patch);
  return merge_value(t, root(t), p, root(p));
}

std::string apply_json_patch(const std::string& target, const std::string& patch, bool* ok) {
  std::string cur = target;
  Doc p = doc_of(patch);
  *ok = true;
  ArrIter it(p, root(p));
  Span e;
  while (it.next(p, e)) {
    const std::string opname = unquote(p, find(p, e, "op"));
    const int op = opname == "add" ? 0 : opname == "remove" ? 1 : opname == "replace" ? 2 : -1;
    const std::string path = unquote(p, find(p, e, "path"));
    if (op < 0 || path.empty() || path[0] != '/') {
      *ok = false;
      return target;
    }
    std::vector<std::string> tokens;
    size_t a = 1;
    for (;;) {
      const size_t b = path.find('/', a);
      tokens.push_back(unescape_pointer(path.substr(a, b == std::string::npos ? b : b - a)));
      if (b == std::string::npos) break;
      a = b + 1;
    }
    Doc d = doc_of(cur);
    std::string next;
    if (!patch_at(d, root(d), tokens, 0, op, text(p, find(p, e, "value")), &next)) {
      *ok = false;
      return target;
    }
    cur = std::move(next);
  }
  return cur;
}

Err Store::Patch(const Key& k, const std::string& merge_patch_json) {
  std::lock_guard<std::mutex> l(mu_);
  auto it = objs_.find(k);
  if (it == objs_.end()) return Err::NotFound;
  std::string o = apply_merge_patch(it->second, merge_patch_json);
  o = with_meta_field(o, "resourceVersion", "\"" + std::to_string(++rv_) + "\"");
  ++cnt_.patches;
  it->second = std::move(o);
  return Err::Ok;
}

Err Store::PatchJson(const Key& k, const std::string& json_patch) {
  std::lock_guard<std::mutex> l(mu_);
  auto it = objs_.find(k);
  if (it == objs_.end()) return Err::NotFound;
  bool ok = false;
  std::string o = apply_json_patch(it->second, json_patch, &ok);
  if (!ok) return Err::Invalid;
  o = with_meta_field(o, "resourceVersion", "\"" + std::to_string(++rv_) + "\"");
  ++cnt_.patches;
  it->second = std::move(o);
  return Err::Ok;
}

Err Store::Get(const Key& k, std::string* out) {
  std::lock_guard<std::mutex> l(mu_);
  ++cnt_.gets;
  auto it = objs_.find(k);
  if (it == objs_.end()) return Err::NotFound;
  if (out) *out = it->second;
  return Err::Ok;
}

Err Store::Create(const std::string& json, std::string* stored) {
  ObjMeta m = read_meta(json);
  if (!m.ok) return Err::Invalid;
  std::lock_guard<std::mutex> l(mu_);
  Key k{m.kind, m.ns, m.name};
  if (objs_.count(k)) return Err::AlreadyExists;
  std::string o = with_meta_field(json, "resourceVersion", "\"" + std::to_string(++rv_) + "\"");
  if (m.uid.empty()) {
    char buf[64];
    snprintf(buf, sizeof buf, "\"00000000-0000-4000-8000-%012llx\"", (unsigned long long)++uid_);
    o = with_meta_field(o, "uid", buf);
  }
  ++cnt_.creates;
  if (stored) *stored = o;
  objs_[k] = std::move(o);
  return Err::Ok;
}

Err Store::Update(const std::string& json, std::string* stored) {
  ObjMeta m = read_meta(json);
  if (!m.ok) return Err::Invalid;
  std::lock_guard<std::mutex> l(mu_);
  Key k{m.kind, m.ns, m.name};
  auto it = objs_.find(k);
  if (it == objs_.end()) return Err::NotFound;
  if (!m.resource_version.empty() && m.resource_version != read_meta(it->second).resource_version) {
    ++cnt_.conflicts;
    return Err::Conflict;
  }
  std::string o = with_meta_field(json, "resourceVersion", "\"" + std::to_string(++rv_) + "\"");
  ++cnt_.updates;
  if (stored) *stored = o;
  it->second = std::move(o);
  return Err::Ok;
}

Err Store::UpdateStatus(const Key& k, const std::string& status_json) {
  std::lock_guard<std::mutex> l(mu_);
  auto it = objs_.find(k);
  if (it == objs_.end()) return Err::NotFound;
  std::string o = with_status(it->second, status_json);
  o = with_meta_field(o, "resourceVersion", "\"" + std::to_string(++rv_) + "\"");
  ++cnt_.status_updates;
  it->second = std::move(o);
  return Err::Ok;
}

Err Store::Delete(const Key& k) {
  std::lock_guard<std::mutex> l(mu_);
  if (!objs_.erase(k)) return Err::NotFound;
  ++cnt_.deletes;
  return Err::Ok;
}

size_t Store::size() const {
  std::lock_guard<std::mutex> l(mu_);
  return objs_.size();
}
Counters Store::counters() const {
  std::lock_guard<std::mutex> l(mu_);
  return cnt_;
}
void Store::reset_counters() {
  std::lock_guard<std::mutex> l(mu_);
  cnt_ = Counters();
}

}  // namespace nbrm_host
