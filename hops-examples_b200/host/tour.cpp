// tour.cpp -- see tour.hpp.  Every compute step below is a call into libfsx (include/fsx.h).
#include "tour.hpp"

#include <dirent.h>
#include <sys/stat.h>

#include <algorithm>
#include <cinttypes>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iterator>
#include <sstream>
#include <stdexcept>

namespace fsxhost {

// ---- small RAII layer over the C ABI ---------------------------------------------------------------
namespace {

struct Fsx {
    fsx_ctx *ctx;
    void ok(int32_t rc, const char *what) const {
        if (rc != FSX_OK) throw std::runtime_error(std::string(what) + ": " + fsx_last_error(ctx));
    }
};

struct Col {
    fsx_ctx *ctx = nullptr;
    fsx_col *h = nullptr;
    bool owned = true;
    Col() = default;
    Col(fsx_ctx *c, fsx_col *p, bool o = true) : ctx(c), h(p), owned(o) {}
    Col(Col &&o) noexcept : ctx(o.ctx), h(o.h), owned(o.owned) { o.h = nullptr; }
    Col &operator=(Col &&o) noexcept { reset(); ctx = o.ctx; h = o.h; owned = o.owned; o.h = nullptr; return *this; }
    Col(const Col &) = delete;
    Col &operator=(const Col &) = delete;
    ~Col() { reset(); }
    void reset() { if (h && owned) fsx_column_free(ctx, h); h = nullptr; }
};

struct Table {
    fsx_ctx *ctx = nullptr;
    fsx_table *h = nullptr;
    Table(fsx_ctx *c, fsx_table *t) : ctx(c), h(t) {}
    Table(const Table &) = delete;
    ~Table() { if (h) fsx_table_free(ctx, h); }
    const fsx_col *col(int i) const { return fsx_table_column(h, i); }
    int64_t nrows() const { return fsx_table_nrows(h); }
};

struct Join {
    fsx_ctx *ctx = nullptr;
    fsx_join *h = nullptr;
    Join(fsx_ctx *c, fsx_join *j) : ctx(c), h(j) {}
    Join(const Join &) = delete;
    ~Join() { if (h) fsx_join_free(ctx, h); }
};

Col upload(fsx_ctx *ctx, const CsvColumn &c) {
    fsx_col *out = nullptr;
    const uint8_t *valid = c.validity.empty() ? nullptr : c.validity.data();
    int32_t rc;
    switch (c.type) {
    case CsvType::Int: rc = fsx_column_upload(ctx, FSX_I32, (int64_t)c.i32.size(), c.i32.data(), valid, &out); break;
    case CsvType::Long: rc = fsx_column_upload(ctx, FSX_I64, (int64_t)c.i64.size(), c.i64.data(), valid, &out); break;
    case CsvType::Double: rc = fsx_column_upload(ctx, FSX_F64, (int64_t)c.f64.size(), c.f64.data(), valid, &out); break;
    default: throw std::runtime_error("column " + c.name + " is a string column; only numeric columns go to the GPU");
    }
    Fsx{ctx}.ok(rc, ("upload " + c.name).c_str());
    return Col(ctx, out);
}

Col upload(fsx_ctx *ctx, const HostColumn &c) {
    fsx_col *out = nullptr;
    Fsx{ctx}.ok(fsx_column_upload(ctx, c.dtype, c.n, c.data.data(), c.validity.empty() ? nullptr : c.validity.data(), &out),
                ("upload " + c.name).c_str());
    return Col(ctx, out);
}

HostColumn download(fsx_ctx *ctx, const std::string &name, const fsx_col *col) {
    HostColumn h;
    h.name = name;
    int32_t has_validity = 0;
    Fsx{ctx}.ok(fsx_column_info(col, &h.dtype, &h.n, &has_validity), "column_info");
    const size_t esz = (h.dtype == FSX_I32 || h.dtype == FSX_F32) ? 4 : 8;
    h.data.resize((size_t)h.n * esz);
    if (has_validity) h.validity.resize((size_t)((h.n + 7) / 8));
    Fsx{ctx}.ok(fsx_column_download(ctx, col, h.data.data(), has_validity ? h.validity.data() : nullptr), ("download " + name).c_str());
    return h;
}

// Spark's typed accessors throw when the inferred type does not match the case class
// (`getAs[Long]("sum(position)")` on a double column, ComputeFeatures.scala:152,210,249-252)
void require_type(const CsvColumn &c, CsvType want, const char *table) {
    if (c.type != want)
        throw std::runtime_error(std::string(table) + ": column '" + c.name + "' was inferred as " + csv_type_name(c.type) +
                                 " but the job reads it as " + csv_type_name(want));
}

std::string fmt_double(double v) {
    if (std::isnan(v)) return "null";
    if (std::isinf(v)) return v > 0 ? "\"Infinity\"" : "\"-Infinity\"";
    char buf[40];
    snprintf(buf, sizeof buf, "%.17g", v);
    return buf;
}

std::string fmt_cell(const HostColumn &c, int64_t i) {
    if (!c.valid(i)) return "";
    char buf[40];
    switch (c.dtype) {
    case FSX_I32: snprintf(buf, sizeof buf, "%d", c.as<int32_t>()[i]); break;
    case FSX_I64: snprintf(buf, sizeof buf, "%" PRId64, c.as<int64_t>()[i]); break;
    case FSX_F32: snprintf(buf, sizeof buf, "%.9g", (double)c.as<float>()[i]); break;
    default: snprintf(buf, sizeof buf, "%.17g", c.as<double>()[i]); break;
    }
    return buf;
}

const char *deequ_type(int32_t dtype) { return (dtype == FSX_I32 || dtype == FSX_I64) ? "Integral" : "Fractional"; }

} // namespace

const HostColumn &DataFrame::col(const std::string &name) const {
    for (const auto &c : cols)
        if (c.name == name) return c;
    throw std::runtime_error("DataFrame column not found: " + name);
}

// ---- DirSink -------------------------------------------------------------------------------------
void DirSink::save(const FeatureGroupMeta &meta, const DataFrame &df, const std::string &statistics_json) {
    mkdir(dir_.c_str(), 0777);
    const std::string base = dir_ + "/" + meta.name + "_" + std::to_string(meta.version);
    FILE *f = fopen((base + ".csv").c_str(), "w");
    if (!f) throw std::runtime_error("cannot write " + base + ".csv");
    for (size_t c = 0; c < df.cols.size(); c++) fprintf(f, "%s%s", c ? "," : "", df.cols[c].name.c_str());
    fputc('\n', f);
    for (int64_t r = 0; r < df.nrows; r++) {
        for (size_t c = 0; c < df.cols.size(); c++) fprintf(f, "%s%s", c ? "," : "", fmt_cell(df.cols[c], r).c_str());
        fputc('\n', f);
    }
    fclose(f);
    f = fopen((base + ".statistics.json").c_str(), "w");
    if (!f) throw std::runtime_error("cannot write " + base + ".statistics.json");
    fputs(statistics_json.c_str(), f);
    fclose(f);
    f = fopen((base + ".meta.json").c_str(), "w");
    if (f) {
        fprintf(f, "{\"name\":\"%s\",\"version\":%d,\"description\":\"%s\",\"timeTravelFormat\":\"%s\",\"onlineEnabled\":%s,\"primaryKeys\":[",
                meta.name.c_str(), meta.version, meta.description.c_str(), meta.time_travel_format.c_str(), meta.online_enabled ? "true" : "false");
        for (size_t i = 0; i < meta.primary_keys.size(); i++) fprintf(f, "%s\"%s\"", i ? "," : "", meta.primary_keys[i].c_str());
        fputs("],\"partitionKeys\":[", f);
        for (size_t i = 0; i < meta.partition_keys.size(); i++) fprintf(f, "%s\"%s\"", i ? "," : "", meta.partition_keys[i].c_str());
        fputs("]}\n", f);
        fclose(f);
    }
    saved_[meta.name] = df;
}

const DataFrame &DirSink::get(const std::string &name) const {
    auto it = saved_.find(name);
    if (it == saved_.end()) throw std::runtime_error("feature group not found: " + name);
    return it->second;
}

// ---- statistics on save (StatisticsConfig(true,true,true), CF:114) ----------------------------------
namespace {
std::string json_escape(const std::string &in) {
    std::string o;
    for (unsigned char c : in) {
        if (c == '"' || c == '\\') { o += '\\'; o += (char)c; }
        else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += (char)c;
    }
    return o;
}
} // namespace

std::string ComputeFeatures::computeStatistics(const std::vector<std::pair<std::string, const fsx_col *>> &all_cols,
                                               bool histograms, bool correlations,
                                               const std::map<std::string, std::vector<std::string>> &dictionaries) {
    const Fsx api{ctx_};
    // numeric columns go through fsx_stats_profile together (shared describe pass, one correlation matrix);
    // dictionary-coded string columns are profiled one by one below
    std::vector<std::pair<std::string, const fsx_col *>> cols, cats;
    for (const auto &c : all_cols) (dictionaries.count(c.first) ? cats : cols).push_back(c);
    const int d = (int)cols.size();
    std::vector<const fsx_col *> handles;
    std::vector<int32_t> dtypes((size_t)d);
    std::vector<int32_t> has_nulls((size_t)d);
    int64_t n = 0;
    for (int i = 0; i < d; i++) {
        handles.push_back(cols[(size_t)i].second);
        api.ok(fsx_column_info(handles.back(), &dtypes[(size_t)i], &n, &has_nulls[(size_t)i]), "column_info");
    }
    std::vector<fsx_moments> mom((size_t)d);
    const int nbins = 20;
    std::vector<int64_t> counts((size_t)d * nbins, 0);
    std::vector<double> edges((size_t)d * (nbins + 1), 0.0);
    std::vector<int32_t> nbuckets((size_t)d, 0);
    std::vector<double> corr;
    bool corr_ok = false;
    if (correlations && d >= 2 && n > 1) {                // Pearson needs null-free columns and more than one row
        bool any_nulls = false;
        for (int i = 0; i < d; i++) any_nulls |= has_nulls[(size_t)i] != 0;
        if (!any_nulls) { corr.resize((size_t)d * d); corr_ok = true; }
    }
    // one library call: describe once, histogram bounds and Gram scaling reuse it
    if (d > 0)
        api.ok(fsx_stats_profile(ctx_, handles.data(), d, nbins, mom.data(), histograms ? counts.data() : nullptr,
                                 histograms ? edges.data() : nullptr, histograms ? nbuckets.data() : nullptr,
                                 corr_ok ? corr.data() : nullptr), "fsx_stats_profile");
    std::vector<char> hist_ok((size_t)d, 0);
    for (int i = 0; i < d; i++) hist_ok[(size_t)i] = histograms && nbuckets[(size_t)i] > 0;
    // value-frequency metrics: Deequ's approximateNumDistinctValues is a HyperLogLog estimate (fsx_stats_approx_distinct);
    // the exact count is reported next to it.  The 100 percentiles of Deequ's ApproxQuantiles are exact order statistics.
    std::vector<fsx_distinct_stats> freq((size_t)d);
    std::vector<int64_t> approx((size_t)d, 0);
    std::vector<double> pct((size_t)d * 100, NAN);
    for (int i = 0; i < d; i++) {
        api.ok(fsx_stats_distinct(ctx_, handles[(size_t)i], &freq[(size_t)i]), "fsx_stats_distinct");
        api.ok(fsx_stats_approx_distinct(ctx_, handles[(size_t)i], &approx[(size_t)i], nullptr), "fsx_stats_approx_distinct");
        for (int first = 0; first < 100; first += 16) {
            double probs[16];
            const int cnt = 100 - first < 16 ? 100 - first : 16;
            for (int k = 0; k < cnt; k++) probs[k] = (first + k + 1) / 100.0;
            api.ok(fsx_stats_quantiles(ctx_, handles[(size_t)i], probs, cnt, &pct[(size_t)i * 100 + first], nullptr), "fsx_stats_quantiles");
        }
    }
    std::ostringstream js;
    js << "{\"columns\":[";
    for (int i = 0; i < d; i++) {
        const fsx_moments &m = mom[(size_t)i];
        const double nn = (double)m.n_nonnull;
        js << (i ? "," : "") << "{\"column\":\"" << json_escape(cols[(size_t)i].first) << "\",\"dataType\":\"" << deequ_type(dtypes[(size_t)i])
           << "\",\"isDataTypeInferred\":\"false\",\"count\":" << m.n_rows
           << ",\"completeness\":" << fmt_double(m.n_rows ? nn / (double)m.n_rows : 1.0)
           << ",\"approximateNumDistinctValues\":" << approx[(size_t)i]
           << ",\"numDistinctValuesExact\":" << freq[(size_t)i].n_distinct
           << ",\"distinctness\":" << fmt_double(m.n_nonnull ? (double)freq[(size_t)i].n_distinct / nn : 0.0)
           << ",\"uniqueness\":" << fmt_double(m.n_nonnull ? (double)freq[(size_t)i].n_unique / nn : 0.0)
           << ",\"entropy\":" << fmt_double(freq[(size_t)i].entropy)
           << ",\"mean\":" << fmt_double(m.n_nonnull ? m.mean : NAN) << ",\"maximum\":" << fmt_double(m.n_nonnull ? m.max : NAN)
           << ",\"minimum\":" << fmt_double(m.n_nonnull ? m.min : NAN) << ",\"sum\":" << fmt_double(m.n_nonnull ? m.sum : NAN)
           << ",\"stdDev\":" << fmt_double(m.n_nonnull ? std::sqrt(m.m2 / nn) : NAN)                      // Deequ: population
           << ",\"stdDevSample\":" << fmt_double(m.n_nonnull > 1 ? std::sqrt(m.m2 / (nn - 1.0)) : NAN);   // Spark describe()
        js << ",\"approxPercentiles\":[";
        if (m.n_nonnull)
            for (int k = 0; k < 100; k++) js << (k ? "," : "") << fmt_double(pct[(size_t)i * 100 + k]);
        js << "]";
        if (histograms) {
            js << ",\"histogram\":[";
            if (hist_ok[(size_t)i]) {
                const int nb = nbuckets[(size_t)i];
                for (int b = 0; b < nb; b++) {
                    const int64_t c = counts[(size_t)i * nbins + b];
                    js << (b ? "," : "") << "{\"value\":" << fmt_double(edges[(size_t)i * (nbins + 1) + b]) << ",\"upper\":"
                       << fmt_double(nb == 1 ? edges[(size_t)i * (nbins + 1)] : edges[(size_t)i * (nbins + 1) + b + 1])
                       << ",\"count\":" << c << ",\"ratio\":" << fmt_double(m.n_nonnull ? (double)c / nn : 0.0) << "}";
                }
            }
            js << "]";
        }
        if (correlations) {
            js << ",\"correlations\":[";
            if (corr_ok)
                for (int j = 0; j < d; j++)
                    js << (j ? "," : "") << "{\"column\":\"" << json_escape(cols[(size_t)j].first) << "\",\"correlation\":"
                       << fmt_double(corr[(size_t)i * d + j]) << "}";
            js << "]";
        }
        js << "}";
    }
    // categorical (string) columns: Deequ's profile of a String column -- completeness, approximate distinct count, and
    // the histogram value -> (count, ratio of all rows, nulls as "NullValue"), most frequent value first
    for (size_t ci = 0; ci < cats.size(); ci++) {
        const fsx_col *col = cats[ci].second;
        const std::vector<std::string> &dict = dictionaries.at(cats[ci].first);
        int64_t rows = 0;
        api.ok(fsx_column_info(col, nullptr, &rows, nullptr), "column_info");
        fsx_distinct_stats fr;
        int64_t ap = 0;
        api.ok(fsx_stats_distinct(ctx_, col, &fr), "fsx_stats_distinct");
        api.ok(fsx_stats_approx_distinct(ctx_, col, &ap, nullptr), "fsx_stats_approx_distinct");
        js << ((d > 0 || ci > 0) ? "," : "") << "{\"column\":\"" << json_escape(cats[ci].first) << "\",\"dataType\":\"String\",\"isDataTypeInferred\":\"false\",\"count\":" << rows
           << ",\"completeness\":" << fmt_double(rows ? (double)fr.n_nonnull / (double)rows : 1.0)
           << ",\"approximateNumDistinctValues\":" << ap << ",\"numDistinctValuesExact\":" << fr.n_distinct
           << ",\"distinctness\":" << fmt_double(fr.n_nonnull ? (double)fr.n_distinct / (double)fr.n_nonnull : 0.0)
           << ",\"uniqueness\":" << fmt_double(fr.n_nonnull ? (double)fr.n_unique / (double)fr.n_nonnull : 0.0)
           << ",\"entropy\":" << fmt_double(fr.entropy);
        if (histograms) {
            fsx_table *vc = nullptr;
            api.ok(fsx_stats_value_counts(ctx_, col, &vc), "fsx_stats_value_counts");
            Table guard(ctx_, vc);
            HostColumn codes = download(ctx_, "code", guard.col(0)), cnt = download(ctx_, "count", guard.col(1));
            js << ",\"histogram\":[";
            for (int64_t r = 0; r < guard.nrows(); r++) {
                const int32_t code = codes.as<int32_t>()[r];
                const std::string value = codes.valid(r) && code >= 0 && (size_t)code < dict.size() ? dict[(size_t)code] : std::string("NullValue");
                const int64_t c = cnt.as<int64_t>()[r];
                js << (r ? "," : "") << "{\"value\":\"" << json_escape(value) << "\",\"count\":" << c << ",\"ratio\":" << fmt_double(rows ? (double)c / (double)rows : 0.0) << "}";
            }
            js << "]";
        }
        js << "}";
    }
    js << "]}\n";
    return js.str();
}

// Every column of a headered CSV: numeric columns as inferred (Spark 2.4 rules, csv.hpp), string columns dictionary-coded
// on the host (fsx_dictionary_encode_utf8) with empty cells as nulls.
std::string ComputeFeatures::profileCsv(const std::string &path, bool histograms, bool correlations) {
    const Fsx api{ctx_};
    CsvTable raw = read_csv(path);
    std::vector<Col> owned;
    std::vector<std::pair<std::string, const fsx_col *>> cols;
    std::map<std::string, std::vector<std::string>> dictionaries;
    for (const CsvColumn &c : raw.cols) {
        if (c.type != CsvType::String) { owned.push_back(upload(ctx_, c)); cols.push_back({c.name, owned.back().h}); continue; }
        const int64_t n = (int64_t)c.raw.size();
        std::vector<int64_t> offsets((size_t)n + 1, 0);
        std::string bytes;
        std::vector<uint8_t> valid((size_t)(n + 7) / 8 + 1, 0);
        bool any_null = false;
        for (int64_t i = 0; i < n; i++) {
            bytes += c.raw[(size_t)i];
            offsets[(size_t)i + 1] = (int64_t)bytes.size();
            if (!c.raw[(size_t)i].empty()) valid[(size_t)(i >> 3)] |= (uint8_t)(1u << (i & 7)); else any_null = true;
        }
        std::vector<int32_t> codes((size_t)n + 1), first((size_t)n + 1);
        int32_t nd = 0;
        api.ok(fsx_dictionary_encode_utf8((const uint8_t *)bytes.data(), offsets.data(), valid.data(), n, codes.data(), first.data(), &nd), "fsx_dictionary_encode_utf8");
        std::vector<std::string> dict;
        for (int32_t k = 0; k < nd; k++) dict.push_back(c.raw[(size_t)first[(size_t)k]]);
        fsx_col *h = nullptr;
        api.ok(fsx_column_upload(ctx_, FSX_I32, n, codes.data(), any_null ? valid.data() : nullptr, &h), ("upload " + c.name).c_str());
        owned.emplace_back(ctx_, h);
        cols.push_back({c.name, owned.back().h});
        dictionaries[c.name] = std::move(dict);
    }
    return computeStatistics(cols, histograms, correlations, dictionaries);
}

// The data files a Delta table consists of now: the JSON commits of <dir>/_delta_log in version order, every "add" entering
// and every "remove" leaving the set (Delta protocol, reader version 1; checkpoints are not read -- the tables the notebooks
// bundle have a handful of commits).  Paths are relative to the table directory and percent-encoded.
std::vector<std::string> deltaActiveFiles(const std::string &dir) {
    const std::string logdir = dir + "/_delta_log";
    DIR *d = opendir(logdir.c_str());
    if (!d) throw std::runtime_error("not a Delta table (no _delta_log): " + dir);
    std::vector<std::string> commits;
    while (dirent *e = readdir(d)) {
        const std::string name = e->d_name;
        if (name.size() == 25 && name.substr(20) == ".json" && name.find_first_not_of("0123456789") == 20) commits.push_back(name);
    }
    closedir(d);
    std::sort(commits.begin(), commits.end());
    auto path_of = [](const std::string &line, size_t from) {                      // the value of "path":"..." after `from`
        const size_t k = line.find("\"path\":\"", from);
        std::string out;
        if (k == std::string::npos) return out;
        for (size_t i = k + 8; i < line.size() && line[i] != '"'; i++) {
            char c = line[i];
            if (c == '\\' && i + 1 < line.size()) c = line[++i];                   // \" \\ \/
            if (c == '%' && i + 2 < line.size() && isxdigit((unsigned char)line[i + 1]) && isxdigit((unsigned char)line[i + 2])) {
                c = (char)strtol(line.substr(i + 1, 2).c_str(), nullptr, 16);
                i += 2;
            }
            out += c;
        }
        return out;
    };
    std::vector<std::string> active;
    for (const std::string &commit : commits) {
        std::ifstream f(logdir + "/" + commit);
        for (std::string line; std::getline(f, line);) {
            const bool add = line.compare(0, 7, "{\"add\":") == 0, remove = line.compare(0, 10, "{\"remove\":") == 0;
            if (!add && !remove) continue;
            const std::string p = path_of(line, 0);
            if (p.empty()) continue;
            active.erase(std::remove(active.begin(), active.end(), p), active.end());
            if (add) active.push_back(p);
        }
    }
    for (std::string &p : active) p = dir + "/" + p;
    return active;
}

// Every decodable column of a table stored as Parquet -- one file, or the part files of a directory in order -- decoded on
// the GPU with all pages of all files in one launch per column (fsx_parquet_decode_column_files / _strings_files).
std::string ComputeFeatures::profileParquet(const std::vector<std::string> &paths, std::vector<std::string> *skipped, bool histograms, bool correlations) {
    const Fsx api{ctx_};
    if (paths.empty()) throw std::runtime_error("profileParquet: no file");
    std::vector<std::string> images;
    for (const std::string &path : paths) {
        std::ifstream f(path, std::ios::binary);
        if (!f) throw std::runtime_error("cannot open " + path);
        images.emplace_back((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    }
    std::vector<const void *> files;
    std::vector<int64_t> sizes;
    for (const std::string &im : images) { files.push_back(im.data()); sizes.push_back((int64_t)im.size()); }
    const int32_t nf = (int32_t)files.size();
    std::vector<char> schema(1 << 16);
    int64_t nrows = 0;
    api.ok(fsx_parquet_schema(files[0], sizes[0], schema.data(), schema.size(), &nrows), "fsx_parquet_schema");
    std::vector<Col> owned;
    std::vector<std::pair<std::string, const fsx_col *>> cols;
    std::map<std::string, std::vector<std::string>> dictionaries;
    std::stringstream items(schema.data());
    for (std::string item; std::getline(items, item, ',');) {
        const size_t colon = item.rfind(':');
        if (colon == std::string::npos) continue;
        const std::string name = item.substr(0, colon);
        std::string type = item.substr(colon + 1);
        if (!type.empty() && type.back() == '?') type.pop_back();
        fsx_col *h = nullptr;
        int32_t rc;
        if (type == "byte_array") {
            int32_t nd = 0;
            int64_t need = 0;
            rc = fsx_parquet_decode_strings_files(ctx_, files.data(), sizes.data(), nf, name.c_str(), &h, nullptr, 0, nullptr, 0, &nd, &need);
            if (rc == FSX_E_INVALID && nd > 0) {                                   // sizes known now
                std::vector<char> bytes((size_t)need + 1);
                std::vector<int64_t> offsets((size_t)nd + 1);
                rc = fsx_parquet_decode_strings_files(ctx_, files.data(), sizes.data(), nf, name.c_str(), &h, bytes.data(), need, offsets.data(), nd + 1, &nd, &need);
                if (rc == FSX_OK) {
                    std::vector<std::string> dict;
                    for (int32_t k = 0; k < nd; k++) dict.emplace_back(bytes.data() + offsets[(size_t)k], (size_t)(offsets[(size_t)k + 1] - offsets[(size_t)k]));
                    dictionaries[name] = std::move(dict);
                }
            } else if (rc == FSX_OK) dictionaries[name] = {};
        } else {
            rc = fsx_parquet_decode_column_files(ctx_, files.data(), sizes.data(), nf, name.c_str(), &h);
        }
        if (rc == FSX_E_UNSUPPORTED) { if (skipped) skipped->push_back(name); continue; }
        api.ok(rc, ("parquet column " + name).c_str());
        owned.emplace_back(ctx_, h);
        cols.push_back({name, owned.back().h});
    }
    return computeStatistics(cols, histograms, correlations, dictionaries);
}

namespace {

// groupBy("team_id").sum() + .count() + join + map for one (sum column list): shared by the three
// aggregating feature groups; returns the saved DataFrame in case-class field order.
struct AggSpec { std::string source; CsvType type; std::string avg_name; std::string sum_name; };

DataFrame aggregate_by_team(ComputeFeatures &self, fsx_ctx *ctx, const CsvTable &raw, const char *table,
                            const std::vector<AggSpec> &specs, const std::vector<std::string> &field_order,
                            std::string *stats_json) {
    const Fsx api{ctx};
    const CsvColumn &team = raw.col("team_id");
    require_type(team, CsvType::Int, table);                                  // row.getAs[Int]("team_id")
    Col key = upload(ctx, team);
    std::vector<Col> vals;
    std::vector<const fsx_col *> val_handles;
    std::vector<int32_t> aggs;
    for (const auto &s : specs) {
        const CsvColumn &c = raw.col(s.source);
        require_type(c, s.type, table);                                       // getAs[Long] / getAs[Double]
        vals.push_back(upload(ctx, c));
        val_handles.push_back(vals.back().h);
        aggs.push_back(FSX_AGG_SUM);
    }
    val_handles.push_back(nullptr);
    aggs.push_back(FSX_AGG_COUNT_STAR);                                       // groupBy(...).count()
    fsx_table *t = nullptr;
    api.ok(fsx_groupby_agg(ctx, key.h, val_handles.data(), aggs.data(), (int32_t)aggs.size(), 0, &t), "fsx_groupby_agg");
    Table tab(ctx, t);
    // `sum.join(count, "team_id")` is an inner join: the null-key group never matches and is dropped.
    // fsx_table_sort_by puts it last; it is cut off below.
    api.ok(fsx_table_sort_by(ctx, tab.h, 0), "fsx_table_sort_by");
    const fsx_col *count = tab.col((int)specs.size() + 1);
    std::map<std::string, Col> feats;
    for (size_t i = 0; i < specs.size(); i++) {
        fsx_col *avg = nullptr, *sum = nullptr;
        api.ok(fsx_finalize_f32(ctx, tab.col((int)i + 1), count, &avg), "fsx_finalize_f32(avg)");
        feats.emplace(specs[i].avg_name, Col(ctx, avg));
        api.ok(fsx_finalize_f32(ctx, tab.col((int)i + 1), nullptr, &sum), "fsx_finalize_f32(sum)");
        feats.emplace(specs[i].sum_name, Col(ctx, sum));
    }
    int64_t G = tab.nrows();
    {   // drop the null-key group if there is one
        int32_t dt, hv;
        int64_t n;
        api.ok(fsx_column_info(tab.col(0), &dt, &n, &hv), "column_info");
        if (hv) {
            int64_t nulls = 0;
            api.ok(fsx_column_null_count(ctx, tab.col(0), &nulls), "null_count");
            G -= nulls;
        }
    }
    std::vector<std::pair<std::string, const fsx_col *>> stat_cols;
    std::vector<Col> views;
    DataFrame df;
    df.nrows = G;
    for (const auto &name : field_order) {
        const fsx_col *src = name == "team_id" ? tab.col(0) : feats.at(name).h;
        fsx_col *view = nullptr;
        api.ok(fsx_column_slice(ctx, src, 0, G, &view), "slice");
        views.emplace_back(ctx, view);
        HostColumn hc = download(ctx, name, view);
        hc.validity.clear();                                                  // after the cut no nulls remain
        df.cols.push_back(std::move(hc));
        stat_cols.emplace_back(name, view);
    }
    // the slices of the key column still carry the (now all-valid) bitmap; statistics see plain views
    std::vector<Col> plain;
    for (auto &sc : stat_cols) {
        void *data = nullptr, *valid = nullptr;
        int32_t dt, hv;
        int64_t n;
        api.ok(fsx_column_device_ptr(sc.second, &data, &valid), "device_ptr");
        api.ok(fsx_column_info(sc.second, &dt, &n, &hv), "column_info");
        fsx_col *v = nullptr;
        api.ok(fsx_column_wrap_device(ctx, dt, n, data, nullptr, &v), "wrap");
        plain.emplace_back(ctx, v);
        sc.second = v;
    }
    *stats_json = self.computeStatistics(stat_cols, true, true);
    return df;
}

FeatureGroupMeta meta_of(const char *name, const char *description, std::vector<std::string> pk) {
    FeatureGroupMeta m;
    m.name = name;
    m.version = ComputeFeatures::FEATUREGROUP_VERSION;
    m.description = description;
    m.primary_keys = std::move(pk);
    return m;
}

} // namespace

// CF:101-132
void ComputeFeatures::computeGamesFeatureGroup(const Logger &log, const std::string &datasetPath) {
    log.info(std::string("Computing feature group: ") + GAMES_FEATUREGROUP);
    const std::string input = datasetPath + SLASH_DELIMITER + GAMES_DATASET_FILE;
    CsvTable raw = read_csv(input);
    // rawDf.as[RawGame]: away_team_id, home_team_id, score -- all Int (CF:86-89); no compute, but save => statistics
    DataFrame df;
    df.nrows = raw.nrows;
    std::vector<Col> dev;
    std::vector<std::pair<std::string, const fsx_col *>> stat_cols;
    for (const char *name : {"away_team_id", "home_team_id", "score"}) {
        const CsvColumn &c = raw.col(name);
        require_type(c, CsvType::Int, "games.csv");
        dev.push_back(upload(ctx_, c));
        df.cols.push_back(download(ctx_, name, dev.back().h));
        stat_cols.emplace_back(name, dev.back().h);
    }
    const std::string stats = computeStatistics(stat_cols, true, true);
    log.info(std::string("Creating featuregroup ") + GAMES_FEATUREGROUP + " version " + std::to_string(FEATUREGROUP_VERSION));
    fs_->save(meta_of(GAMES_FEATUREGROUP, "Features of games", {"home_team_id"}), df, stats);
    log.info(std::string("Creation of featuregroup ") + GAMES_FEATUREGROUP + " complete");
    FeatureGroupMeta hudi = meta_of(GAMES_FEATUREGROUP_TOUR_HUDI, "Features of games, HUDI feature group example", {"home_team_id"});
    hudi.time_travel_format = "HUDI";
    hudi.partition_keys = {"score"};
    fs_->save(hudi, df, stats);                                               // hudi_fg.save(rawDs.toDF), CF:131
    log.info(std::string("Creation Hudi featuregroup ") + GAMES_FEATUREGROUP_TOUR_HUDI + " complete");
}

// CF:142-177
void ComputeFeatures::computeSeasonScoresFeatureGroup(const Logger &log, const std::string &datasetPath) {
    log.info(std::string("Computing feature group: ") + SEASON_SCORES_FEATUREGROUP);
    CsvTable raw = read_csv(datasetPath + SLASH_DELIMITER + SEASON_SCORES_DATASET_FILE);
    std::string stats;
    DataFrame df = aggregate_by_team(*this, ctx_, raw, "season_scores.csv",
                                     {{"position", CsvType::Int, "average_position", "sum_position"}},
                                     {"team_id", "average_position", "sum_position"}, &stats);    // SeasonScoreTeamFeature, CF:81-84
    FeatureGroupMeta m = meta_of(SEASON_SCORES_FEATUREGROUP, "Features of average season scores for football teams", {"team_id"});
    m.online_enabled = true;                                                   // CF:168
    fs_->save(m, df, stats);
    log.info(std::string("Creation of featuregroup ") + SEASON_SCORES_FEATUREGROUP + " complete");
}

// CF:179-191: an on-demand feature group is a stored SQL query against the online store
// ("SELECT * FROM season_scores_features_1 WHERE average_position > 3"); it has no compute in the job
// and its storage connector is out of scope, so only the definition is recorded.
void ComputeFeatures::computeSeasonScoresOnDemandFeatureGroup(const Logger &log) {
    FeatureGroupMeta m = meta_of(SEASON_FEATUREGROUP_TOUR_ON_DEMAND, "Features of games, on demand feature group example", {});
    DataFrame empty;
    fs_->save(m, empty, "{\"query\":\"SELECT * FROM season_scores_features_1 WHERE average_position > 3\"}\n");
    log.info(std::string("Creation of on-demand featuregroup ") + SEASON_FEATUREGROUP_TOUR_ON_DEMAND + " complete");
}

// CF:200-230
void ComputeFeatures::computeAttendanceFeatureGroup(const Logger &log, const std::string &datasetPath) {
    log.info(std::string("Computing feature group: ") + ATTENDANCES_FEATUREGROUP);
    CsvTable raw = read_csv(datasetPath + SLASH_DELIMITER + ATTENDANCES_DATASET_FILE);
    std::string stats;
    DataFrame df = aggregate_by_team(*this, ctx_, raw, "attendances.csv",
                                     {{"attendance", CsvType::Double, "average_attendance", "sum_attendance"}},
                                     {"team_id", "average_attendance", "sum_attendance"}, &stats); // AttendanceTeamFeature, CF:71-74
    fs_->save(meta_of(ATTENDANCES_FEATUREGROUP, "Features of average attendance of games of football teams", {"team_id"}), df, stats);
    log.info(std::string("Creation of featuregroup ") + ATTENDANCES_FEATUREGROUP + " complete");
}

// CF:239-277
void ComputeFeatures::computePlayersFeatureGroup(const Logger &log, const std::string &datasetPath) {
    log.info(std::string("Computing feature group: ") + PLAYERS_FEATUREGROUP);
    CsvTable raw = read_csv(datasetPath + SLASH_DELIMITER + PLAYERS_DATASET_FILE);
    std::string stats;
    DataFrame df = aggregate_by_team(*this, ctx_, raw, "players.csv",
                                     {{"age", CsvType::Int, "average_player_age", "sum_player_age"},
                                      {"rating", CsvType::Double, "average_player_rating", "sum_player_rating"},
                                      {"worth", CsvType::Double, "average_player_worth", "sum_player_worth"}},
                                     {"team_id", "average_player_rating", "average_player_age", "average_player_worth",
                                      "sum_player_rating", "sum_player_age", "sum_player_worth"}, &stats); // PlayersTeamFeature, CF:57-64
    fs_->save(meta_of(PLAYERS_FEATUREGROUP, "Aggregate features of players football teams", {"team_id"}), df, stats);
    log.info(std::string("Creation of featuregroup ") + PLAYERS_FEATUREGROUP + " complete");
}

// CF:286-306
void ComputeFeatures::computeTeamsFeatureGroup(const Logger &log, const std::string &datasetPath) {
    log.info(std::string("Computing feature group: ") + TEAMS_FEATUREGROUP);
    CsvTable raw = read_csv(datasetPath + SLASH_DELIMITER + TEAMS_DATASET_FILE);
    const Fsx api{ctx_};
    const CsvColumn &budget = raw.col("team_budget"), &id = raw.col("team_id"), &pos = raw.col("team_position");
    require_type(budget, CsvType::Double, "teams.csv");
    require_type(id, CsvType::Int, "teams.csv");
    require_type(pos, CsvType::Int, "teams.csv");
    Col b = upload(ctx_, budget), i = upload(ctx_, id), p = upload(ctx_, pos);
    fsx_col *bf = nullptr;
    api.ok(fsx_finalize_f32(ctx_, b.h, nullptr, &bf), "team_budget.toFloat");     // CF:292
    Col budget_f(ctx_, bf);
    DataFrame df;                                                                  // TeamFeature(team_budget, team_id, team_position), CF:45-48
    df.nrows = raw.nrows;
    df.cols.push_back(download(ctx_, "team_budget", budget_f.h));
    df.cols.push_back(download(ctx_, "team_id", i.h));
    df.cols.push_back(download(ctx_, "team_position", p.h));
    const std::string stats = computeStatistics({{"team_budget", budget_f.h}, {"team_id", i.h}, {"team_position", p.h}}, true, true);
    fs_->save(meta_of(TEAMS_FEATUREGROUP, "Features of football teams", {"team_id"}), df, stats);
    log.info(std::string("Creation of featuregroup ") + TEAMS_FEATUREGROUP + " complete");
}

// CF:312-329: players_fg.select(average_player_age) JOIN teams_fg.select(team_budget) JOIN
// attendance_fg.select(average_attendance) -- hsfs joins on the primary key team_id, INNER.
void ComputeFeatures::createTrainingDataset(const Logger &log) {
    log.info(std::string("Creating Training Dataset: ") + TOUR_TRAINING_DATASET);
    const Fsx api{ctx_};
    const DataFrame &players = fs_->get(PLAYERS_FEATUREGROUP), &teams = fs_->get(TEAMS_FEATUREGROUP),
                    &att = fs_->get(ATTENDANCES_FEATUREGROUP);
    Col pk = upload(ctx_, players.col("team_id")), tk = upload(ctx_, teams.col("team_id")), ak = upload(ctx_, att.col("team_id"));
    Col age = upload(ctx_, players.col("average_player_age")), budget = upload(ctx_, teams.col("team_budget")),
        attendance = upload(ctx_, att.col("average_attendance"));
    fsx_join *j1h = nullptr, *j2h = nullptr;
    api.ok(fsx_hash_join(ctx_, pk.h, tk.h, FSX_JOIN_INNER, &j1h), "players JOIN teams");
    Join j1(ctx_, j1h);
    auto gather = [&](const Join &j, int side, const fsx_col *c) {
        fsx_col *o = nullptr;
        api.ok(fsx_join_gather(ctx_, j.h, side, c, &o), "fsx_join_gather");
        return Col(ctx_, o);
    };
    Col k1 = gather(j1, 0, pk.h), age1 = gather(j1, 0, age.h), budget1 = gather(j1, 1, budget.h);
    api.ok(fsx_hash_join(ctx_, k1.h, ak.h, FSX_JOIN_INNER, &j2h), "(players JOIN teams) JOIN attendances");
    Join j2(ctx_, j2h);
    Col age2 = gather(j2, 0, age1.h), budget2 = gather(j2, 0, budget1.h), att2 = gather(j2, 1, attendance.h);
    DataFrame df;
    df.cols.push_back(download(ctx_, "average_player_age", age2.h));
    df.cols.push_back(download(ctx_, "team_budget", budget2.h));
    df.cols.push_back(download(ctx_, "average_attendance", att2.h));
    df.nrows = df.cols[0].n;
    const std::string stats = computeStatistics({{"average_player_age", age2.h}, {"team_budget", budget2.h},
                                                 {"average_attendance", att2.h}}, true, true);
    FeatureGroupMeta m = meta_of(TOUR_TRAINING_DATASET, "Sample Training Dataset for the Feature store Tour", {});
    m.time_travel_format = "TFRECORD";                                           // dataFormat(DataFormat.TFRECORD), CF:324
    fs_->save(m, df, stats);
}

// Main.scala:60-66
std::string parseInput(const std::string &input, const std::string &projectName) {
    if (input.find("hdfs://") != std::string::npos) return input;
    if (!input.empty() && (input[0] == '/' || input[0] == '.')) return input;     // local runs: keep file-system paths
    return "hdfs:///Projects/" + projectName + "/" + input;
}

// Main.scala:42-48
void runTourJob(fsx_ctx *ctx, const std::string &input, FeatureGroupSink *sink, const Logger &log) {
    log.info("Starting Sample Feature Engineering Job For Feature Store Examples");
    if (input.find("hdfs://") != std::string::npos)
        throw std::runtime_error("this build reads the local file system only; got " + input);
    ComputeFeatures cf(ctx, sink);
    cf.computeGamesFeatureGroup(log, input);
    cf.computeSeasonScoresFeatureGroup(log, input);
    cf.computeAttendanceFeatureGroup(log, input);
    cf.computePlayersFeatureGroup(log, input);
    cf.computeTeamsFeatureGroup(log, input);
    cf.computeSeasonScoresOnDemandFeatureGroup(log);
    cf.createTrainingDataset(log);
    log.info("Shutting down spark job");
}

} // namespace fsxhost
