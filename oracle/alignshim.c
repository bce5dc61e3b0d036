/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * The unmodified reference wraps a std::malloc'ed arena with
 * ggml_backend_cpu_buffer_from_ptr (acestep_ggml/cpp/acestep_dit_model.cpp:794,
 * acestep_vae_model.cpp:829), which asserts 32-byte alignment
 * (third_party/ggml/src/ggml-backend.cpp:2265).  glibc malloc only guarantees 16,
 * so the reference aborts on Linux.  This LD_PRELOAD shim makes every malloc
 * 64-byte aligned; no reference source is edited.
 */
#include <stddef.h>
extern void *__libc_memalign(size_t, size_t);
void *malloc(size_t n) { return __libc_memalign(64, n ? n : 1); }
