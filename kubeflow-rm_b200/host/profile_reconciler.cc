// profile_reconciler.cc — see profile_reconciler.h.
#include "profile_reconciler.h"

#include <map>

#include "../csrc/nbrm_cmp.cuh"

namespace nbrm_host {

namespace {
using nbrm::Doc;
using nbrm::Span;

Doc doc_of(const std::string& s) {
  Doc d;
  d.p = s.data();
  d.n = (uint32_t)s.size();
  return d;
}

// Raw JSON text of a nested member ("" when absent).
std::string raw_at(const std::string& json, const std::vector<std::string>& path) {
  const Doc d = doc_of(json);
  Span cur = nbrm::root(d);
  for (const std::string& key : path) {
    Span next = nbrm::mk(0, 0);
    nbrm::ObjIter it(d, cur);
    nbrm::Member m;
    while (it.next(d, m))
      if (std::string(d.p + m.key.off + 1, m.key.len - 2) == key) next = m.val;
    cur = next;
  }
  return std::string(d.p + cur.off, cur.len);
}

bool is_nullish(const std::string& raw) { return raw.empty() || raw == "null"; }

// reflect.DeepEqual of two decoded values, on their JSON (absent and null are the same nil).
bool same_json(const std::string& a, const std::string& b) {
  if (is_nullish(a) || is_nullish(b)) return is_nullish(a) == is_nullish(b);
  const Doc da = doc_of(a), db = doc_of(b);
  return nbrm::json_equal(da, nbrm::root(da), db, nbrm::root(db));
}

std::string quote(const std::string& s) {
  std::string o = "\"";
  for (unsigned char c : s) {
    if (c == '"' || c == '\\') {
      o += '\\';
      o += (char)c;
    } else if (c < 0x20) {
      char b[8];
      snprintf(b, sizeof b, "\\u%04x", c);
      o += b;
    } else {
      o += (char)c;
    }
  }
  return o + "\"";
}

// The elements of a JSON array as raw texts.
std::vector<std::string> elements(const std::string& raw) {
  std::vector<std::string> out;
  const Doc d = doc_of(raw);
  nbrm::ArrIter it(d, nbrm::root(d));
  Span e;
  while (it.next(d, e)) out.push_back(std::string(d.p + e.off, e.len));
  return out;
}

std::string array_of(const std::vector<std::string>& items) {
  std::string o = "[";
  for (size_t i = 0; i < items.size(); ++i) o += (i ? "," : "") + items[i];
  return o + "]";
}

// A JSON object of strings as key -> raw value (keys decoded as they stand between the quotes).
std::map<std::string, std::string> string_map(const std::string& raw) {
  std::map<std::string, std::string> m;
  const Doc d = doc_of(raw);
  nbrm::ObjIter it(d, nbrm::root(d));
  nbrm::Member mem;
  while (it.next(d, mem))
    m[std::string(d.p + mem.key.off, mem.key.len)] = std::string(d.p + mem.val.off, mem.val.len);
  return m;
}

std::string object_of(const std::map<std::string, std::string>& m) {
  std::string o = "{";
  bool first = true;
  for (auto& kv : m) {
    o += (first ? "" : ",") + kv.first + ":" + kv.second;
    first = false;
  }
  return o + "}";
}

// `json` with the top-level or nested member at `pointer` set to the JSON value `raw`.
std::string set_at(const std::string& json, const std::string& pointer, const std::string& raw, bool* ok) {
  return apply_json_patch(json, "[{\"op\":\"add\",\"path\":\"" + pointer + "\",\"value\":" + raw + "}]", ok);
}

struct Plugin {
  std::string kind, spec;
};

std::vector<Plugin> plugins_of(const std::string& profile) {  // GetPluginSpec :676-714
  std::vector<Plugin> out;
  for (const std::string& p : elements(raw_at(profile, {"spec", "plugins"}))) {
    const std::string kind = read_string(p, {"kind"});
    if (kind != "WorkloadIdentity" && kind != "AwsIamForServiceAccount") continue;  // "Plugin not recgonized"
    out.push_back(Plugin{kind, raw_at(p, {"spec"})});
  }
  return out;
}
}  // namespace

bool ProfileReconciler::ReconcileOne(std::string profile, const std::string objs[8], ProfileStats& st,
                                     std::string* error) {
  const std::string name = read_string(profile, {"metadata", "name"});
  const std::string owner = read_string(profile, {"spec", "owner", "name"});
  auto fail = [&](const char* what) {
    *error = what;
    return false;
  };
  auto write_profile = [&](const std::string& next) {  // r.Update(ctx, instance): the instance takes the stored form
    std::string stored;
    if (store_->Update(next, &stored) != Err::Ok) return false;
    profile = stored;
    ++st.profile_writes;
    return true;
  };
  // appendErrorConditionAndReturn :337-347
  auto append_failed = [&](const std::string& message) {
    std::vector<std::string> conds = elements(raw_at(profile, {"status", "conditions"}));
    conds.push_back("{\"type\":\"Failed\",\"message\":" + quote(message) + "}");
    bool ok = true;
    std::string next = is_nullish(raw_at(profile, {"status"}))
                           ? set_at(profile, "/status", "{\"conditions\":" + array_of(conds) + "}", &ok)
                           : set_at(profile, "/status/conditions", array_of(conds), &ok);
    return ok && write_profile(next);
  };

  // ---- Namespace :127-204
  {
    std::string found;
    const Err e = store_->Get(Key{"Namespace", "", name}, &found);
    if (e == Err::NotFound) {
      if (store_->Create(objs[0]) != Err::Ok) return fail("error creating namespace");
      ++st.creates;
      if (store_->Get(Key{"Namespace", "", name}, &found) != Err::Ok)  // the 5 x 3 s wait :160-165
        return append_failed("Owning namespace failed to create within 15 seconds") || fail("error updating the profile");
    } else if (e != Err::Ok) {
      return fail("error reading namespace");
    } else {
      const std::string ann = read_annotation(found, "owner");
      if (!ann.empty() && read_string(found, {"metadata", "annotations", "owner"}) == owner) {  // :176-191
        const std::map<std::string, std::string> old = string_map(raw_at(found, {"metadata", "labels"}));
        std::map<std::string, std::string> now = old;
        for (auto& kv : string_map(labels_)) {  // setNamespaceLabels :754-772
          if (kv.second == "\"\"") now.erase(kv.first);
          else if (!now.count(kv.first)) now[kv.first] = kv.second;
        }
        if (now != old) {
          if (store_->Update(with_meta_field(found, "labels", object_of(now))) != Err::Ok)
            return fail("error updating namespace label");
          ++st.updates;
        }
      } else {  // :192-203: the namespace belongs to someone else
        ++st.rejected;
        return append_failed("namespace already exist, but not owned by profile creator " + owner) ||
               fail("error updating the profile");
      }
    }
  }
  // ---- objects whose reconcilers own some fields: create, or set those fields when they differ
  auto create_or_update = [&](const std::string& desired, const char* kind, const std::vector<std::string>& owned) {
    const ObjMeta m = read_meta(desired);
    std::string found;
    const Err e = store_->Get(Key{kind, m.ns, m.name}, &found);
    if (e == Err::NotFound) {
      if (store_->Create(desired) != Err::Ok) return false;
      ++st.creates;
      return true;
    }
    if (e != Err::Ok) return false;
    bool differs = false;
    for (const std::string& f : owned)
      if (!same_json(raw_at(desired, {f}), raw_at(found, {f}))) differs = true;
    if (!differs) return true;
    bool ok = true;
    for (const std::string& f : owned) {  // found.<field> = desired.<field>, every owned field (:580,:664-665)
      const std::string v = raw_at(desired, {f});
      found = set_at(found, "/" + f, v.empty() ? "null" : v, &ok);
      if (!ok) return false;
    }
    if (store_->Update(found) != Err::Ok) return false;
    ++st.updates;
    return true;
  };
  if (!create_or_update(objs[1], "AuthorizationPolicy", {"spec"}))  // updateIstioAuthorizationPolicy :509-558
    return fail("error updating Istio AuthorizationPolicy permission");
  for (int k : {2, 4}) {  // updateServiceAccount :592-637: the account is only ever created
    if (!create_or_update(objs[k], "ServiceAccount", {})) return fail("error updating ServiceAccount");
    if (!create_or_update(objs[k + 1], "RoleBinding", {"roleRef", "subjects"})) return fail("error updating ServiceAccount");
  }
  if (!create_or_update(objs[6], "RoleBinding", {"roleRef", "subjects"}))  // the owner's binding :211-246
    return fail("error updating Owner Rolebinding");
  if (!objs[7].empty()) {  // :253-262
    if (!create_or_update(objs[7], "ResourceQuota", {"spec"})) return fail("error updating resource quota");
  } else {  // :263-277: a quota the Profile no longer asks for goes away
    std::string found;
    const Err e = store_->Get(Key{"ResourceQuota", name, "kf-resource-quota"}, &found);
    if (e == Err::Ok) {
      if (store_->Delete(Key{"ResourceQuota", name, "kf-resource-quota"}) != Err::Ok) return fail("error deleting resource quota");
      ++st.deletes;
    } else if (e != Err::NotFound) {
      return fail("error getting resource quota");
    }
  }
  // ---- PatchDefaultPluginSpec :716-740: the default plugin if its kind is not there, then Update — always
  {
    std::string next = profile;
    if (!wi_.empty()) {
      std::vector<std::string> plugins = elements(raw_at(profile, {"spec", "plugins"}));
      bool have = false;
      for (const std::string& p : plugins)
        if (read_string(p, {"kind"}) == "WorkloadIdentity") have = true;
      if (!have) {
        plugins.push_back("{\"kind\":\"WorkloadIdentity\",\"spec\":{\"gcpServiceAccount\":" + quote(wi_) + "}}");
        bool ok = true;
        next = set_at(profile, "/spec/plugins", array_of(plugins), &ok);
        if (!ok) return fail("error patching DefaultPluginSpec");
      }
    }
    if (!write_profile(next)) return fail("error patching DefaultPluginSpec");
  }
  for (const Plugin& p : plugins_of(profile)) {  // :285-291
    if (apply_ && !apply_(p.kind, p.spec, profile)) return fail("error applying plugin");
    ++st.plugins_applied;
  }
  // ---- the finalizer :292-330
  {
    std::vector<std::string> fins = elements(raw_at(profile, {"metadata", "finalizers"}));
    bool have = false;
    for (const std::string& f : fins)
      if (f == "\"profile-finalizer\"") have = true;
    if (!read_meta(profile).has_deletion_timestamp) {
      if (!have) {
        fins.push_back("\"profile-finalizer\"");
        if (!write_profile(with_meta_field(profile, "finalizers", array_of(fins)))) return fail("error updating finalizer");
        ++st.finalizers_added;
      }
    } else if (have) {
      for (const Plugin& p : plugins_of(profile)) {
        if (revoke_ && !revoke_(p.kind, p.spec, profile)) return fail("error revoking plugin");
        ++st.plugins_revoked;
      }
      std::vector<std::string> rest;
      for (const std::string& f : fins)
        if (f != "\"profile-finalizer\"") rest.push_back(f);
      if (!write_profile(with_meta_field(profile, "finalizers", rest.empty() ? "" : array_of(rest))))
        return fail("error removing finalizer");
      ++st.finalizers_removed;
    }
  }
  return true;
}

int ProfileReconciler::ReconcileBatch(const std::vector<std::string>& names, ProfileStats* stats,
                                      std::vector<std::string>* errors) {
  ProfileStats st;
  st.keys = names.size();
  std::vector<std::string> errs(names.size());
  std::vector<size_t> live;
  std::vector<std::string> profiles;
  for (size_t i = 0; i < names.size(); ++i) {
    std::string p;
    if (store_->Get(Key{"Profile", "", names[i]}, &p) != Err::Ok) {  // :112-120
      ++st.not_found;
      continue;
    }
    live.push_back(i);
    profiles.push_back(std::move(p));
  }
  if (!live.empty()) {
    std::string data;
    std::vector<uint64_t> off{0};
    for (auto& p : profiles) {
      data += p;
      off.push_back(data.size());
    }
    nbrm_docs d;
    d.data = data.data();
    d.off = off.data();
    d.n = profiles.size();
    nbrm_out out;
    if (nbrm_profile_generate(ctx_, &d, &out) < 0) {
      for (size_t j = 0; j < live.size(); ++j) errs[live[j]] = nbrm_last_error(ctx_);
    } else {
      // keep the objects: the context's host buffers are reused by the next call
      std::vector<std::string> objs(live.size() * 8);
      std::vector<uint32_t> gst(out.status, out.status + live.size());
      for (size_t o = 0; o < objs.size(); ++o) objs[o].assign(out.data + out.off[o], out.off[o + 1] - out.off[o]);
      for (size_t j = 0; j < live.size(); ++j) {
        if (gst[j]) {
          errs[live[j]] = "the Profile cannot be handled";
          continue;
        }
        if (ReconcileOne(profiles[j], &objs[8 * j], st, &errs[live[j]])) ++st.reconciled;
      }
    }
  }
  for (auto& e : errs)
    if (!e.empty()) ++st.errors;
  if (stats) *stats = st;
  if (errors) *errors = std::move(errs);
  return 0;
}

}  // namespace nbrm_host
