"""Test helper (CPU only): a mutation fuzzer for the Parquet decoder of csrc/parquet.cu.

The source between `namespace {` and the device driver -- the Thrift footer / page-header walk that ships as host code and
the three kernels, compiled for the host with a one-lane warp exactly as tests/pq_emulation.py does -- is built into a
stand-alone executable with AddressSanitizer and UBSan.  Its main() keeps every scratch area at the size decode_on_device
allocates on the GPU (so a read or write the device code would make outside its allocations is reported here), mutates
valid files (byte flips, small integers, truncation, spliced ranges) and decodes every column of every mutant.  A mutant
may be rejected with a status or decode to other values; it must never crash, hang or touch memory outside the buffers.

Run by hand for long sessions:  python tests/pq_fuzz.py 200000
The CPU suite runs a bounded number of mutants (tests/test_parquet_cpu.py).
"""
import os
import subprocess
import sys

import pq_emulation

MAIN = r'''
#include <cstdlib>
#include <fstream>
#include <iterator>
static uint64_t rng_state = 88172645463325252ULL;
static uint64_t rnd() { rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17; return rng_state; }

// one column of one (possibly corrupt) file through the plan and the three kernels; buffers sized like decode_on_device's
static int decode_one(const std::vector<uint8_t> &file, const std::string &column, bool as_string, long long max_rows) {
    PqPlan plan; PqStrings strings; std::string err;
    int type = -1, optional = 0;
    const void *fp = file.data();
    const int64_t fn = (int64_t)file.size();
    const int rc = plan_column(&fp, &fn, 1, column.c_str(), plan, type, optional, as_string ? &strings : nullptr, err);
    if (rc) return rc;
    const int elem = as_string ? 4 : pq_elem(type);
    const long long nrows = plan.nrows;
    if (nrows > max_rows || plan.bufpos > (64LL << 20)) return -100;       // a header that claims gigabytes: allocation is the caller's limit
    std::vector<uint8_t> image((size_t)plan.image_len + 16), buf((size_t)plan.bufpos + 64), lv((size_t)nrows + 64);
    std::vector<uint8_t> ixb((size_t)nrows * 4 + 64);
    std::vector<uint8_t> out((size_t)nrows * pq_out_elem(elem) + 0), valid((size_t)((nrows + 127) / 128) * 16);
    for (auto &r : plan.ranges) memcpy(image.data() + r.image_pos, r.file + r.file_pos, (size_t)r.len);
    for (auto &u : strings.uploads) if (!u.codes.empty()) memcpy(buf.data() + u.buf_pos, u.codes.data(), u.codes.size() * 4);
    for (auto &hd : plan.host_dicts) {
        std::vector<uint8_t> plain;
        const uint8_t *bytes = hd.src;
        if (hd.codec == 1) { if (!snappy_decode_host(hd.src, (size_t)hd.compressed, (size_t)hd.uncompressed, plain)) return FSX_E_DATA; bytes = plain.data(); }
        if (hd.uncompressed > 0) memcpy(buf.data() + hd.buf_pos, bytes, (size_t)hd.uncompressed);
    }
    unsigned bad = 0;
    const std::vector<PqPageDev> &pages = plan.pages;
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_snappy(image.data(), buf.data(), pages.data(), &bad); }
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_decode(buf.data(), pages.data(), optional, elem, lv.data(), (unsigned *)ixb.data(), &bad); }
    for (size_t i = 0; i < pages.size(); i++) { blockIdx.x = (int)i; k_pq_place(buf.data(), pages.data(), optional, elem, lv.data(), (const unsigned *)ixb.data(), out.data(), (unsigned *)valid.data(), &bad); }
    return bad == 2 ? FSX_E_UNSUPPORTED : bad ? FSX_E_DATA : 0;
}

int main(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: pq_fuzz <mutants> <seed> <file>...\n"); return 2; }
    const long mutants = atol(argv[1]);
    rng_state ^= (uint64_t)atoll(argv[2]) * 0x9E3779B97F4A7C15ULL;
    std::vector<std::vector<uint8_t>> seeds;
    std::vector<std::vector<std::pair<std::string, bool>>> columns;
    for (int i = 3; i < argc; i++) {
        std::ifstream f(argv[i], std::ios::binary);
        std::vector<uint8_t> bytes((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
        PqFile pf; std::string err;
        if (!parse_footer(bytes.data(), (int64_t)bytes.size(), pf, err)) { fprintf(stderr, "%s: %s\n", argv[i], err.c_str()); return 2; }
        std::vector<std::pair<std::string, bool>> cols;
        for (auto &c : pf.cols) cols.push_back({c.name, c.type == PQ_BYTE_ARRAY});
        long ok = 0;
        for (auto &c : cols) ok += decode_one(bytes, c.first, c.second, 1LL << 22) == 0;
        fprintf(stderr, "seed %s: %zu bytes, %zu columns, %ld decode cleanly\n", argv[i], bytes.size(), cols.size(), ok);
        seeds.push_back(bytes); columns.push_back(cols);
    }
    long by_status[8] = {0};
    for (long m = 0; m < mutants; m++) {
        const size_t s = rnd() % seeds.size();
        std::vector<uint8_t> f = seeds[s];
        const int edits = 1 + (int)(rnd() % 4);
        for (int e = 0; e < edits; e++) {
            const unsigned kind = (unsigned)(rnd() % 16);
            // a third of the edits go to the footer (the last ~2 KB), where every byte is structure
            size_t at = rnd() % f.size();
            if (kind & 1) { const size_t tail = f.size() < 2048 ? f.size() : 2048; at = f.size() - 1 - rnd() % tail; }
            if (kind < 6) f[at] ^= (uint8_t)(1u << (rnd() % 8));
            else if (kind < 10) f[at] = (uint8_t)rnd();
            else if (kind < 12) f[at] = (uint8_t)((rnd() & 1) ? 0xFF : 0x00);
            else if (kind == 12) { const size_t n = rnd() % 16; for (size_t k = 0; k < n && at + k < f.size(); k++) f[at + k] = (uint8_t)rnd(); }
            else if (kind == 13 && f.size() > 64) { const size_t from = rnd() % f.size(), n = rnd() % 64; for (size_t k = 0; k < n && at + k < f.size() && from + k < f.size(); k++) f[at + k] = f[from + k]; }
            else if (kind == 14 && f.size() > 32) { const size_t cut = rnd() % 24; f.resize(f.size() - cut); }                    // truncation (also removes the magic)
            else if (f.size() > 32) { const size_t n = 1 + rnd() % 8; f.erase(f.begin() + (long)at, f.begin() + (long)std::min(f.size(), at + n)); }
        }
        for (auto &c : columns[s]) {
            const int rc = decode_one(f, c.first, c.second, 1LL << 22);
            by_status[rc == 0 ? 0 : rc == FSX_E_INVALID ? 1 : rc == FSX_E_UNSUPPORTED ? 2 : rc == FSX_E_DATA ? 3 : rc == -100 ? 4 : 5]++;
        }
    }
    printf("decodes: %ld ok, %ld invalid, %ld unsupported, %ld corrupt data, %ld oversized, %ld other\n", by_status[0], by_status[1], by_status[2], by_status[3],
           by_status[4], by_status[5]);
    return by_status[5] ? 1 : 0;
}
'''


def build(workdir: str, sanitize: bool = True) -> str:
    src = open(pq_emulation.SRC).read()
    a, b = src.index("namespace {"), src.index("// ---- device driver")
    body = src[a:b] + "\n} // namespace\n"
    body = body.replace("constexpr int kLanes = 32;", "constexpr int kLanes = 1;").replace("constexpr int kThreads = 8 * kLanes;", "constexpr int kThreads = 1;")
    cpp = os.path.join(workdir, "pq_fuzz.cpp")
    with open(cpp, "w") as f:
        f.write(pq_emulation.PRELUDE + body + MAIN)
    exe = os.path.join(workdir, "pq_fuzz")
    flags = ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer"] if sanitize else []
    subprocess.run(["g++", "-O1", "-g", "-std=c++17", *flags, "-o", exe, cpp], check=True)
    return exe


def seed_files(workdir: str):
    """small valid files: the golden Delta parts and files written with Spark's settings (nulls, several row groups, small pages)"""
    import parquet_cases as cases
    golden = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "g8_telco_delta")
    paths = [os.path.join(golden, f) for f in sorted(os.listdir(golden)) if f.endswith(".parquet")]
    variants = [(700, 0.2, "snappy", True, 300, 512), (500, 0.5, "snappy", False, None, 1024), (300, 0.0, "none", True, 100, 256), (900, 0.1, "snappy", True, None, 2048)]
    for i, (n, nullp, codec, dictionary, rg, page) in enumerate(variants):
        nullable, required = cases.tables(n, nullp)
        for j, tab in enumerate((nullable, required, cases.spark_types_table(n, nullp))):
            kw = {"use_deprecated_int96_timestamps": True} if j == 2 else {}
            p = os.path.join(workdir, f"seed_{i}_{j}.parquet")
            with open(p, "wb") as f:
                f.write(cases.write(tab, codec, dictionary, rg, page, **kw))
            paths.append(p)
    return paths


def run(exe: str, paths, mutants: int, seed: int = 1, timeout: int = 3600):
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
    return subprocess.run([exe, str(mutants), str(seed), *paths], capture_output=True, text=True, timeout=timeout, env=env)


if __name__ == "__main__":
    import tempfile
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    with tempfile.TemporaryDirectory() as d:
        r = run(build(d), seed_files(d), n, seed)
        sys.stderr.write(r.stderr[-6000:])
        print(r.stdout, "exit", r.returncode)
