/* link_both.c — a C program that links libnbrm and libnbrm_host together, the way
 * INTEGRATION.md tells an embedder to: every release function must free what its own
 * library handed out.  `store` runs the host-library half (no GPU needed); `all` adds the
 * pinned allocation of libnbrm. */
#include <stdio.h>
#include <string.h>

#include "../../include/nbrm.h"
#include "../../include/nbrm_host.h"

int main(int argc, char** argv) {
  const char* obj = "{\"apiVersion\":\"v1\",\"kind\":\"ConfigMap\",\"metadata\":{\"name\":\"a\",\"namespace\":\"n\"}}";
  nbrm_store* s = nbrm_store_new();
  if (!s || nbrm_store_create(s, obj, strlen(obj)) != 0) return 1;
  for (int i = 0; i < 1000; ++i) {
    char* out = NULL;
    uint64_t len = 0;
    if (nbrm_store_get(s, "ConfigMap", "n", "a", &out, &len) != 0 || !out || len < strlen(obj)) return 2;
    nbrm_store_free_string(out); /* malloc'ed by libnbrm_host */
  }
  nbrm_store_free(s);
  if (argc > 1 && !strcmp(argv[1], "all")) {
    for (int i = 0; i < 16; ++i) {
      void* p = nbrm_host_alloc(1 << 20); /* pinned, from libnbrm */
      if (!p) return 3;
      memset(p, i, 1 << 20);
      nbrm_host_free(p);
    }
  }
  puts("ok");
  return 0;
}
