// nbrm_pool.h — host-side construction of the literal pool (fixed template
// fragments, then the JSON-escaped environment strings).
#pragma once
#include <stdio.h>
#include <string.h>

#include <string>
#include <utility>
#include <vector>

#include "../../include/nbrm.h"
#include "nbrm_plan.cuh"

namespace nbrm {

inline std::string json_escape(const char* s) {
  std::string o;
  for (const unsigned char* p = (const unsigned char*)s; *p; ++p) {
    unsigned char ch = *p;
    switch (ch) {
      case '"': o += "\\\""; break;
      case '\\': o += "\\\\"; break;
      case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break;
      case '\t': o += "\\t"; break;
      default:
        if (ch < 0x20) {
          char b[8];
          snprintf(b, sizeof b, "\\u%04x", ch);
          o += b;
        } else {
          o += (char)ch;
        }
    }
  }
  return o;
}

// setNamespaceLabels (profile_controller.go:754-772) applied to the Namespace the
// reconciler builds (:127-137: labels {"istio-injection": "enabled"}): a default
// with an empty value removes the key, any other default is added unless the key
// exists.  labels_json: the labels file as a JSON object of strings (may be null).
// Returns the members of the resulting map as JSON text ("k":"v","k2":"v2").
inline std::string render_namespace_labels(const char* labels_json) {
  std::vector<std::pair<std::string, std::string>> m;  // raw JSON strings, with quotes
  m.push_back({"\"istio-injection\"", "\"enabled\""});
  if (labels_json && *labels_json) {
    Doc d;
    d.p = labels_json;
    d.n = (uint32_t)strlen(labels_json);
    ObjIter it(d, root(d));
    Member mem;
    while (it.next(d, mem)) {
      if (!is_string(d, mem.val)) continue;
      const std::string k(d.p + mem.key.off, mem.key.len), v(d.p + mem.val.off, mem.val.len);
      size_t at = m.size();
      for (size_t i = 0; i < m.size(); ++i)
        if (m[i].first == k) at = i;
      if (v == "\"\"") {
        if (at < m.size()) m.erase(m.begin() + at);
      } else if (at == m.size()) {
        m.push_back({k, v});
      }
    }
  }
  std::string out;
  for (size_t i = 0; i < m.size(); ++i) out += (i ? "," : "") + m[i].first + ":" + m[i].second;
  return out;
}

inline void build_pool_host(const nbrm_env* env, std::string& pool, EnvLits& L) {
  pool.assign(L__TOTAL, '\0');
#define X(name, text) memcpy(&pool[L_##name], text, sizeof(text) - 1);
  NBRM_LITS(X)
#undef X
  auto add = [&](const std::string& s) {
    PoolSpan sp;
    sp.off = (uint32_t)pool.size();
    sp.len = (uint32_t)s.size();
    pool += s;
    pool += '\0';
    return sp;
  };
  auto get = [](const char* v, const char* dflt) { return (v && *v) ? v : dflt; };
  // os.Getenv + len()==0 test: unset and empty both take the default
  // (notebook_controller.go:548-551,:557-560)
  L.istio_host = add("\"" + json_escape(get(env ? env->istio_host : nullptr, "*")) + "\"");
  L.istio_gateway =
      add("\"" + json_escape(get(env ? env->istio_gateway : nullptr, "kubeflow/kubeflow-gateway")) +
          "\"");
  // os.LookupEnv: set-but-empty is kept (:537-539)
  L.cluster_domain =
      add(json_escape((env && env->cluster_domain) ? env->cluster_domain : "cluster.local"));
  L.ctrl_namespace = add(json_escape(get(env ? env->controller_namespace : nullptr, "")));
  L.oauth_image = add(json_escape(get(
      env ? env->oauth_proxy_image : nullptr,
      "registry.redhat.io/openshift4/ose-oauth-proxy@sha256:"
      "4f8d66597feeb32bb18699326029f9a71a5aca4a57679d636b876377c2e95695")));
  // profile-controller strings (GetEnvDefault: unset or empty takes the default)
  L.prof_when_key = add("\"request.headers[" +
                        json_escape(get(env ? env->userid_header : nullptr, "x-goog-authenticated-user-email")) + "]\"");
  L.prof_prefix = add(json_escape(get(env ? env->userid_prefix : nullptr, "accounts.google.com:")));
  L.prof_igw = add("\"" + json_escape(get(env ? env->istio_ingress_gateway_principal : nullptr,
                       "cluster.local/ns/istio-system/sa/istio-ingressgateway-service-account")) + "\"");
  L.prof_kfp = add("\"" + json_escape(get(env ? env->kfp_ui_principal : nullptr,
                       "cluster.local/ns/kubeflow/sa/ml-pipeline-ui")) + "\"");
  L.prof_nbc = add("\"" + json_escape(get(env ? env->notebook_controller_principal : nullptr,
                       "cluster.local/ns/kubeflow/sa/notebook-controller-service-account")) + "\"");
  L.prof_ns_labels = add(render_namespace_labels(env ? env->namespace_labels_json : nullptr));
  // tensorboard-controller: os.LookupEnv, a variable that is set but empty is used as it is
  // (tensorboard_controller.go:535-541); RWO_PVC_SCHEDULING :471-482
  auto quoted = [&](const char* v) { return add("\"" + json_escape(v ? v : "") + "\""); };
  L.tb_image = quoted(env ? env->tensorboard_image : nullptr);
  L.tb_gateway = quoted(env ? env->istio_gateway : nullptr);
  L.tb_host = quoted(env ? env->istio_host : nullptr);
  L.tb_env = 0;
  if (env && env->tensorboard_image) L.tb_env |= 1u;
  if (env && env->istio_gateway) L.tb_env |= 2u;
  if (env && env->istio_host) L.tb_env |= 4u;
  if (env && env->rwo_pvc_scheduling) {
    const char* r = env->rwo_pvc_scheduling;
    if (!strcmp(r, "true") || !strcmp(r, "True") || !strcmp(r, "TRUE")) L.tb_env |= 8u;
    else if (strcmp(r, "false") && strcmp(r, "False") && strcmp(r, "FALSE")) L.tb_env |= 16u;
  }
  // pvcviewer webhook: DEFAULT_POD_SPEC_PATH's PodSpec as JSON; the members other than `volumes`
  // and the items of `volumes` go to the pool separately because Default appends a volume
  // (pvcviewer_webhook.go:135-143)
  {
    std::string rest, vols;
    bool set = false;
    const char* js = env ? env->pvcviewer_default_podspec_json : nullptr;
    if (js && *js) {
      Doc d;
      d.p = js;
      d.n = (uint32_t)strlen(js);
      ObjIter it(d, root(d));
      Member mem;
      while (it.next(d, mem)) {
        if (!is_null(d, mem.val)) set = true;
        if (str_eq(d, mem.key, "volumes")) {
          if (is_array(d, mem.val) && mem.val.len > 2) vols.assign(d.p + mem.val.off + 1, mem.val.len - 2);
        } else {
          rest += (rest.empty() ? "" : ",") + std::string(d.p + mem.all.off, mem.all.len);
        }
      }
      if (!set) {  // DeepEqual(defaultPodSpec, PodSpec{}): the built-in default is used
        rest.clear();
        vols.clear();
      }
    }
    L.pv_default_rest = add(rest);
    L.pv_default_volumes = add(vols);
    L.pv_default_set = set ? 1u : 0u;
  }
  // ADD_FSGROUP: add the fsGroup when unset or exactly "true" (:475)
  L.add_fsgroup = (!env || !env->add_fsgroup || strcmp(env->add_fsgroup, "true") == 0) ? 1u : 0u;
  L.use_istio = (env && env->use_istio && strcmp(env->use_istio, "true") == 0) ? 1u : 0u;
}

}  // namespace nbrm
