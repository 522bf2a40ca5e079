// Test helper: dlopen a libbert (this repo's or the compiled reference's), run bert_params_parse on the remaining arguments
// and print the parsed fields -- tests/test_host.py compares the two libraries line for line.
#include <dlfcn.h>
#include <cstdio>
#include "bert.h"

int main(int argc, char ** argv) {
    if (argc < 2) return 2;
    void * h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    using parse_t = bool (*)(int, char **, bert_params &);
    parse_t parse = (parse_t) dlsym(h, "bert_params_parse");
    if (!parse) return 2;
    bert_params p;
    argv[1] = argv[0];                                   // the library sees: prog args...
    const bool ok = parse(argc - 1, argv + 1, p);
    printf("ok=%d n_threads=%d port=%d model=%s prompt=%s\n", (int) ok, p.n_threads, p.port, p.model, p.prompt);
    return 0;
}
