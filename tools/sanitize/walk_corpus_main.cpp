#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
extern "C" int emul_walk_compare(const uint8_t* data, size_t size, int threads, int* used_pool, int* status);
int main(int argc, char** argv) {
    int diffs = 0;
    for (int a = 1; a < argc; ++a) {
        FILE* f = fopen(argv[a], "rb"); fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
        uint8_t* d = (uint8_t*)malloc(n ? n : 1); if (n) { size_t r = fread(d, 1, n, f); (void)r; } fclose(f);
        for (int th = 0; th < 2; ++th) { int u, s; int rc = emul_walk_compare(d, n, th, &u, &s); if (rc) { ++diffs; printf("%s threads=%d rc=%d st=%d\n", argv[a], th, rc, s); } }
        free(d);
    }
    printf("files=%d diffs=%d\n", argc - 1, diffs); return diffs != 0;
}
