// store.cc — see store.h.
#include "store.h"

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
