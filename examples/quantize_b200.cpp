// quantize_b200 -- same command line as the reference's models/quantize.cpp:271-300
//   usage: quantize_b200 model-f32-or-f16.bin model-quant.bin type      (type 2 = q4_0, 3 = q4_1)
// The block quantisation runs on the GPU (bertx_quantize_file) and writes the reference tool's bytes.
#include <cstdio>
#include <cstdlib>

#include "bert_b200.h"

int main(int argc, char ** argv) {
    if (argc != 4) {
        fprintf(stderr, "usage: %s model-f32.bin model-quant.bin type\n", argv[0]);
        fprintf(stderr, "  type = 2 - q4_0\n");
        fprintf(stderr, "  type = 3 - q4_1\n");
        return 1;
    }
    const int rc = bertx_quantize_file(argv[1], argv[2], atoi(argv[3]));
    if (rc != 0) fprintf(stderr, "%s: failed to quantize model from '%s'\n", argv[0], argv[1]);
    return rc == 0 ? 0 : 1;
}
