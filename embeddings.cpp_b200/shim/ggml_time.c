/* libggml.so for consumers of this libbert.so: the five timer helpers the reference's examples call
 * (ggml/include/ggml/ggml.h:566-570) on CLOCK_MONOTONIC.  Nothing else of ggml exists here. */
#include "ggml.h"

#include <time.h>

static int64_t now_ns(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (int64_t) ts.tv_sec * 1000000000ll + ts.tv_nsec;
}

void    ggml_time_init(void)     {}
int64_t ggml_time_ms(void)       { return now_ns() / 1000000; }
int64_t ggml_time_us(void)       { return now_ns() / 1000; }
int64_t ggml_cycles(void)        { return now_ns(); }          /* 1 "cycle" = 1 ns */
int64_t ggml_cycles_per_ms(void) { return 1000000; }
