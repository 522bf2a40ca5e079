// convert_b200 -- the reference's models/convert-to-ggml.py command line (convert-to-ggml.py:10-51) on this repo's converter:
//   usage: convert_b200 dir-model [ftype]        ftype 0 = f32, 1 = f16 (default), and, beyond the script, 2 = q4_0, 3 = q4_1
// Reads dir-model/{model.safetensors, config.json, vocab.txt} and writes dir-model/ggml-model-{f32,f16,q4_0,q4_1}.bin --
// for 0 / 1 the script's bytes (host only), for 2 / 3 the bytes of script + `quantize` (models/run_conversions.sh; the
// block quantiser runs on the GPU).
#include <cstdio>
#include <cstdlib>
#include <string>

#include "bert_b200.h"

int main(int argc, char ** argv) {
    if (argc < 2 || argc > 3) {
        fprintf(stderr, "Usage: %s dir-model [ftype]\n", argv[0]);
        fprintf(stderr, "  ftype == 0 -> float32\n  ftype == 1 -> float16 (default)\n  ftype == 2 -> q4_0\n  ftype == 3 -> q4_1\n");
        return 1;
    }
    const std::string dir = argv[1];
    const int ftype = argc > 2 ? atoi(argv[2]) : 1;
    static const char * const names[] = {"f32", "f16", "q4_0", "q4_1"};
    if (ftype < 0 || ftype > 3) {
        fprintf(stderr, "Invalid ftype: %d\n", ftype);        // the script's message (convert-to-ggml.py:48-50)
        return 1;
    }
    const std::string out = dir + "/ggml-model-" + names[ftype] + ".bin";
    const int rc = bertx_convert_safetensors((dir + "/model.safetensors").c_str(), (dir + "/config.json").c_str(),
                                             (dir + "/vocab.txt").c_str(), ftype, out.c_str());
    if (rc != 0) {
        fprintf(stderr, "%s: conversion of '%s' failed (%d)\n", argv[0], dir.c_str(), rc);
        return 1;
    }
    printf("Done. Output file: %s\n", out.c_str());
    return 0;
}
