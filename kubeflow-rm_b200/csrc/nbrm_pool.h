// nbrm_pool.h — host-side construction of the literal pool (fixed template
// fragments, then the JSON-escaped environment strings).
#pragma once
#include <stdio.h>
#include <string.h>

#include <string>

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

inline void build_pool_host(const nbrm_env* env, std::string& pool, EnvLits& L) {
  pool.assign(L__TOTAL, '\0');
#define X(name, text) memcpy(&pool[L_##name], text, sizeof(text) - 1);
  NBRM_LITS(X)
#undef X
  auto add = [&](const std::string& s) {
    Span sp = mk((uint32_t)pool.size(), (uint32_t)s.size());
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
  // ADD_FSGROUP: add the fsGroup when unset or exactly "true" (:475)
  L.add_fsgroup = (!env || !env->add_fsgroup || strcmp(env->add_fsgroup, "true") == 0) ? 1u : 0u;
  L.use_istio = (env && env->use_istio && strcmp(env->use_istio, "true") == 0) ? 1u : 0u;
}

}  // namespace nbrm
