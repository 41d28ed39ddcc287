// stubs.cu -- entry points declared in include/b200hnsw.h whose kernels are not written yet.  They fail loudly.
#include "index.hpp"
using namespace b200;
#define STUB(name)                                  \
  set_error(name ": not implemented in this build"); \
  return B200HNSW_ERR_UNSUPPORTED
extern "C" {
int b200exact_create(int32_t, int32_t, int32_t, int64_t, b200exact_t **) { STUB("exact_create"); }
void b200exact_destroy(b200exact_t *) {}
int64_t b200exact_len(const b200exact_t *) { return 0; }
int b200exact_add_batch(b200exact_t *, const float *, int64_t, int64_t *) { STUB("exact_add_batch"); }
int b200exact_remove(b200exact_t *, const int64_t *, int64_t) { STUB("exact_remove"); }
int b200exact_search_batch(b200exact_t *, const float *, int64_t, int32_t, int64_t *, float *, int32_t *,
                           b200hnsw_stats *) { STUB("exact_search_batch"); }
int b200exact_search_batch_device(b200exact_t *, const float *, int64_t, int32_t, int64_t *, float *, int32_t *,
                                  void *) { STUB("exact_search_batch_device"); }
}
