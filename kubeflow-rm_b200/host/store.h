// store.h — in-memory API object store: the part of client.Client the Notebook
// reconcile path uses, with the API server behaviours that path depends on.
//
// Mirrors the calls NotebookReconciler.Reconcile makes (components/
// notebook-controller/controllers/notebook_controller.go): r.Get :126,:147,:189,
// :221, r.Create :151,:192, r.Update :174,:205,:264, r.Status().Update :287,
// r.Delete :254; errors as apierrs.IsNotFound / IsAlreadyExists / IsConflict
// report them.  Objects are kept as JSON text, keyed by kind/namespace/name;
// resourceVersion is a counter bumped on every write, checked on Update when the
// caller sends one (optimistic concurrency).  It plays the role envtest plays in
// the reference's own tests.
#pragma once
#include <stdint.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace nbrm_host {

enum class Err { Ok = 0, NotFound = 1, AlreadyExists = 2, Conflict = 3, Invalid = 4 };

struct Key {
  std::string kind, ns, name;
  bool operator<(const Key& o) const {
    if (kind != o.kind) return kind < o.kind;
    if (ns != o.ns) return ns < o.ns;
    return name < o.name;
  }
};

struct Counters {
  uint64_t gets = 0, creates = 0, updates = 0, status_updates = 0, deletes = 0, conflicts = 0, patches = 0;
};

// Fields of an object's JSON text the store and the reconciler need.
struct ObjMeta {
  std::string kind, ns, name, resource_version, uid;
  bool has_deletion_timestamp = false;
  bool ok = false;
};
ObjMeta read_meta(const std::string& json);
// A string member of a nested object, e.g. read_string(json, {"involvedObject","kind"}); "" when absent.
std::string read_string(const std::string& json, const std::vector<std::string>& path);
// metadata.annotations[key] as the raw JSON string value (with quotes), or "" when absent.
std::string read_annotation(const std::string& json, const std::string& key);
// The same object with metadata.<field> set to the JSON value `raw` (or removed when raw is empty).
std::string with_meta_field(const std::string& json, const std::string& field, const std::string& raw);
// The same object with its top-level "status" replaced by the JSON value `raw`.
std::string with_status(const std::string& json, const std::string& raw);
// The same object without metadata.annotations[key].
std::string without_annotation(const std::string& json, const std::string& key);

// RFC 7386: `target` with the JSON merge patch applied (null deletes a member, objects merge
// member by member, anything else replaces).  What client.RawPatch(types.MergePatchType, ...)
// makes the API server do (odh-notebook-controller/controllers/notebook_controller.go:141-145).
std::string apply_merge_patch(const std::string& target, const std::string& patch);
// RFC 6902: `target` with the operations of `patch` applied in order (add / remove / replace;
// paths are JSON pointers, "-" appends to an array).  ok is false when a path does not resolve.
std::string apply_json_patch(const std::string& target, const std::string& patch, bool* ok);

class Store {
 public:
  Err Get(const Key& k, std::string* out);
  Err Create(const std::string& json, std::string* stored = nullptr);
  Err Update(const std::string& json, std::string* stored = nullptr);
  Err UpdateStatus(const Key& k, const std::string& status_json);
  // r.Patch(ctx, obj, client.RawPatch(types.MergePatchType, patch)) / (types.JSONPatchType, patch)
  Err Patch(const Key& k, const std::string& merge_patch_json);
  Err PatchJson(const Key& k, const std::string& json_patch);
  Err Delete(const Key& k);
  size_t size() const;
  Counters counters() const;
  void reset_counters();

 private:
  mutable std::mutex mu_;
  std::map<Key, std::string> objs_;
  uint64_t rv_ = 0, uid_ = 0;
  Counters cnt_;
};

}  // namespace nbrm_host
