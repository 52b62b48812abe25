// tour_capi.cpp -- C entry points over the host-side mirror (tour.hpp) so that tests can drive
// the job through ctypes, the same way a JVM would drive it through JNI.
#include <cstring>
#include <string>

#include "tour.hpp"

#include <sys/stat.h>

using namespace fsxhost;

static void put(char *dst, size_t len, const std::string &s) {
    if (!dst || !len) return;
    strncpy(dst, s.c_str(), len - 1);
    dst[len - 1] = 0;
}

extern "C" {

// Main.main(--input <dir>) with a DirSink in place of hsfs persistence.  Returns 0 or -1 (+ message).
int32_t fsx_tour_run(const char *input_dir, const char *output_dir, int32_t device, int32_t quiet, char *err, size_t errlen) {
    fsx_ctx *ctx = nullptr;
    int32_t rc = fsx_ctx_create(&device, 1, &ctx);
    if (rc != FSX_OK) { put(err, errlen, fsx_last_error(nullptr)); return rc; }
    int32_t status = 0;
    try {
        DirSink sink(output_dir);
        Logger log;
        log.quiet = quiet != 0;
        runTourJob(ctx, input_dir, &sink, log);
    } catch (const std::exception &e) {
        put(err, errlen, e.what());
        status = -1;
    }
    fsx_ctx_destroy(ctx);
    return status;
}

// "name:type,name:type,..." as inferred from the file (CPU only; host logic of K0)
int32_t fsx_tour_infer_schema(const char *csv_path, char *out, size_t outlen, int64_t *nrows) {
    try {
        CsvTable t = read_csv(csv_path);
        std::string s;
        for (size_t i = 0; i < t.cols.size(); i++) {
            if (i) s += ",";
            s += t.cols[i].name + ":" + csv_type_name(t.cols[i].type) + (t.cols[i].has_nulls ? "?" : "");
        }
        put(out, outlen, s);
        if (nrows) *nrows = t.nrows;
        return 0;
    } catch (const std::exception &e) {
        put(out, outlen, e.what());
        return -1;
    }
}

// Statistics JSON of every column of a CSV file, string columns included (categorical histogram); *needed receives the
// length of the JSON (call again with a larger buffer when it exceeds outlen).
int32_t fsx_tour_profile_csv(const char *csv_path, int32_t device, char *out, size_t outlen, int64_t *needed) {
    fsx_ctx *ctx = nullptr;
    int32_t rc = fsx_ctx_create(&device, 1, &ctx);
    if (rc != FSX_OK) { put(out, outlen, fsx_last_error(nullptr)); return rc; }
    int32_t status = 0;
    try {
        ComputeFeatures cf(ctx, nullptr);
        const std::string js = cf.profileCsv(csv_path);
        if (needed) *needed = (int64_t)js.size() + 1;
        put(out, outlen, js);
    } catch (const std::exception &e) {
        put(out, outlen, e.what());
        status = -1;
    }
    fsx_ctx_destroy(ctx);
    return status;
}

// The same for a Parquet table, its columns decoded on the GPU (ComputeFeatures::profileParquet): `path` is one .parquet file
// or a Delta table directory (then the part files its _delta_log leaves active are read as one table); skipped (may be NULL)
// receives the comma-separated names of the columns libfsx declined.
int32_t fsx_tour_profile_parquet(const char *path, int32_t device, char *out, size_t outlen, int64_t *needed, char *skipped, size_t skipped_len) {
    fsx_ctx *ctx = nullptr;
    int32_t rc = fsx_ctx_create(&device, 1, &ctx);
    if (rc != FSX_OK) { put(out, outlen, fsx_last_error(nullptr)); return rc; }
    int32_t status = 0;
    try {
        ComputeFeatures cf(ctx, nullptr);
        std::vector<std::string> left_out, paths;
        struct stat st;
        if (stat(path, &st) == 0 && S_ISDIR(st.st_mode)) paths = deltaActiveFiles(path);
        else paths.push_back(path);
        const std::string js = cf.profileParquet(paths, &left_out);
        if (needed) *needed = (int64_t)js.size() + 1;
        put(out, outlen, js);
        std::string names;
        for (const std::string &n : left_out) names += (names.empty() ? "" : ",") + n;
        if (skipped) put(skipped, skipped_len, names);
    } catch (const std::exception &e) {
        put(out, outlen, e.what());
        status = -1;
    }
    fsx_ctx_destroy(ctx);
    return status;
}

// The active part files of a Delta table directory, one path per line (host only; no GPU).
int32_t fsx_tour_delta_files(const char *dir, char *out, size_t outlen) {
    try {
        std::string lines;
        for (const std::string &p : deltaActiveFiles(dir)) lines += p + "\n";
        put(out, outlen, lines);
        return 0;
    } catch (const std::exception &e) {
        put(out, outlen, e.what());
        return -1;
    }
}

// One column of the HOST reader (csv.hpp) as raw 8-byte cells (int / bigint widened to int64, double as is) plus one
// validity byte per row; *type = 1 int, 2 bigint, 3 double, 4 string.  For the GPU decoder's parity tests (CPU only).
int32_t fsx_tour_read_csv_column(const char *csv_path, int32_t col, int64_t cap, uint64_t *cells, uint8_t *valid, int32_t *type, int64_t *nrows) {
    try {
        CsvTable t = read_csv(csv_path);
        if (col < 0 || col >= (int32_t)t.cols.size()) return -1;
        const CsvColumn &c = t.cols[(size_t)col];
        *type = (int32_t)c.type;
        *nrows = t.nrows;
        if (t.nrows > cap) return -2;
        for (int64_t r = 0; r < t.nrows; r++) {
            valid[r] = c.validity.empty() ? 1 : ((c.validity[(size_t)(r >> 3)] >> (r & 7)) & 1);
            uint64_t v = 0;
            if (c.type == CsvType::Int) v = (uint64_t)(int64_t)c.i32[(size_t)r];
            else if (c.type == CsvType::Long) v = (uint64_t)c.i64[(size_t)r];
            else if (c.type == CsvType::Double) memcpy(&v, &c.f64[(size_t)r], 8);
            else valid[r] = !c.raw[(size_t)r].empty();
            cells[r] = v;
        }
        return 0;
    } catch (const std::exception &) {
        return -3;
    }
}

int32_t fsx_tour_parse_input(const char *input, const char *project, char *out, size_t outlen) {
    put(out, outlen, parseInput(input, project));
    return 0;
}

} // extern "C"
