#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
extern "C" int emul_walk_compare(const uint8_t* data, size_t size, int threads, int* used_pool, int* status);
int main(int argc, char** argv) {
    FILE* f = fopen(argv[1], "rb"); fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> d(n); fread(d.data(), 1, n, f); fclose(f);
    int bad = 0;
    for (int i = 0; i < 200; ++i) { int u, s; int rc = emul_walk_compare(d.data(), n, 1, &u, &s); if (rc || s || u != 2) { ++bad; printf("iter %d rc %d st %d used %d\n", i, rc, s, u); } }
    printf("done bad=%d\n", bad); return bad != 0;
}
