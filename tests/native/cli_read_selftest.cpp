// Test harness for the command line tool's file reader (csrc/wordcount_main.cpp: read_at), which the GPU-less
// test run cannot reach through the tool itself: prints bytes [lo, hi) of FILE to stdout, read the way
// feed_range() reads them, through a slab of CAP bytes, buffered (mode 0) or with O_DIRECT (mode 1; exit 3
// if the file system refuses it).
//   cli_read_selftest FILE MODE CAP LO HI
#define main wordcount_cli_main
#include "../../cuda-mapreduce_b200/csrc/wordcount_main.cpp"
#undef main

int main(int argc, char **argv) {
    if (argc != 6) return 2;
    const bool direct = atoi(argv[2]) != 0;
    const size_t cap = strtoull(argv[3], nullptr, 10);
    const uint64_t lo = strtoull(argv[4], nullptr, 10), hi = strtoull(argv[5], nullptr, 10);
    const int fd = open(argv[1], O_RDONLY | (direct ? O_DIRECT : 0));
    if (fd < 0) return 3;
    void *slab = nullptr;
    if (posix_memalign(&slab, kIoAlign, cap) != 0) return 1;
    for (uint64_t pos = lo; pos < hi;) {
        char *at = nullptr;
        const ssize_t k = read_at(fd, direct, (char *)slab, cap, pos, (size_t)(hi - pos < cap ? hi - pos : cap), &at);
        if (k < 0) return direct ? 3 : 1;
        if (k == 0) break;
        fwrite(at, 1, (size_t)k, stdout);
        pos += (uint64_t)k;
    }
    free(slab);
    close(fd);
    return 0;
}
