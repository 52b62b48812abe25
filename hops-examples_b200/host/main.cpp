// main.cpp -- CLI mirror of Main.scala:25-52:  fsx_tour --input <dir> [--output <dir>] [--device N]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

extern "C" int fsx_tour_run(const char *, const char *, int, int, char *, size_t);

int main(int argc, char **argv) {
    std::string input, output = "fsx_tour_out";
    int device = 0;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--input") && i + 1 < argc) input = argv[++i];        // Conf.input, Main.scala:13-16
        else if (!strcmp(argv[i], "--output") && i + 1 < argc) output = argv[++i];
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else { fprintf(stderr, "usage: fsx_tour --input <dir with the five csv files> [--output <dir>] [--device N]\n"); return 2; }
    }
    if (input.empty()) { fprintf(stderr, "--input is required\n"); return 2; }
    char err[1024] = {0};
    int rc = fsx_tour_run(input.c_str(), output.c_str(), device, 0, err, sizeof err);
    if (rc) { fprintf(stderr, "fsx_tour failed: %s\n", err); return 1; }
    return 0;
}
