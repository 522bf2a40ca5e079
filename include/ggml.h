/*
 * ggml.h -- timer shim for the reference's example programs.
 *
 * examples/main.cpp:2,9-10,25,32,36,58,68, examples/test_batch_encode.cpp:2,10-11,26,33,56,60,73,
 * examples/server.cpp:2 and examples/test_tokenizer.cpp:2 include "ggml.h" and link `bert ggml`
 * (examples/CMakeLists.txt:16-29), but the only ggml symbols they use are the wall-clock helpers
 * declared at ggml/include/ggml/ggml.h:566-570.  This engine contains no ggml; the header and
 * libggml.so (embeddings.cpp_b200/shim/ggml_time.c) provide just those helpers so that the example
 * sources compile and link unmodified against this libbert.so.
 */
#ifndef GGML_H
#define GGML_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGML_API __attribute__((visibility("default")))

GGML_API void    ggml_time_init(void);      /* call once at program start */
GGML_API int64_t ggml_time_ms(void);
GGML_API int64_t ggml_time_us(void);
GGML_API int64_t ggml_cycles(void);
GGML_API int64_t ggml_cycles_per_ms(void);

#ifdef __cplusplus
}
#endif

#endif /* GGML_H */
