// csv.hpp -- headered CSV reader with Spark-2.4-style schema inference (host side of K0).
//
// Mirrors what `spark.read.format("csv").option("header", true).option("inferSchema", "true")
// .load(input)` yields for the tour tables (ComputeFeatures.scala:104,145,203,242,289): per column
// the narrowest of Integer -> Long -> Double -> String that fits every non-empty cell; empty cells
// are null and do not affect the type.  Decimal/Timestamp/Boolean inference
// is not needed by any tour table and maps to String.  Parsing stays on the host by design; the
// typed columns go to HBM through fsx_column_upload.
#pragma once
#include <cerrno>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace fsxhost {

enum class CsvType { Null = 0, Int = 1, Long = 2, Double = 3, String = 4 };

inline const char *csv_type_name(CsvType t) {
    switch (t) {
    case CsvType::Null: return "null";
    case CsvType::Int: return "int";
    case CsvType::Long: return "bigint";
    case CsvType::Double: return "double";
    default: return "string";
    }
}

struct CsvColumn {
    std::string name;
    CsvType type = CsvType::Null;
    std::vector<std::string> raw;      // cell text ("" = null)
    std::vector<int32_t> i32;
    std::vector<int64_t> i64;
    std::vector<double> f64;
    std::vector<uint8_t> validity;     // Arrow bitmap, empty when the column has no nulls
    bool has_nulls = false;
};

struct CsvTable {
    std::vector<CsvColumn> cols;
    int64_t nrows = 0;
    const CsvColumn &col(const std::string &name) const {
        for (const auto &c : cols)
            if (c.name == name) return c;
        throw std::runtime_error("CSV column not found: " + name);
    }
};

inline bool parse_long(const std::string &s, long long &out) {
    if (s.empty()) return false;
    errno = 0;
    char *end = nullptr;
    // Java Integer/Long.parseLong: optional sign, digits only, no surrounding blanks
    size_t i = (s[0] == '+' || s[0] == '-') ? 1 : 0;
    if (i == s.size()) return false;
    for (size_t k = i; k < s.size(); k++)
        if (s[k] < '0' || s[k] > '9') return false;
    out = strtoll(s.c_str(), &end, 10);
    return errno == 0 && end == s.c_str() + s.size();
}

inline bool parse_double(const std::string &s, double &out) {
    if (s.empty()) return false;
    // reject forms strtod takes but Java's Double.parseDouble does not (hex floats, "inf", "nan")
    for (char ch : s)
        if (ch == 'x' || ch == 'X') return false;
    if (s == "NaN") { out = std::numeric_limits<double>::quiet_NaN(); return true; }
    if (s == "Infinity" || s == "+Infinity") { out = INFINITY; return true; }
    if (s == "-Infinity") { out = -INFINITY; return true; }
    for (char ch : s)
        if (!((ch >= '0' && ch <= '9') || ch == '+' || ch == '-' || ch == '.' || ch == 'e' || ch == 'E')) return false;
    errno = 0;
    char *end = nullptr;
    out = strtod(s.c_str(), &end);
    return end == s.c_str() + s.size();
}

// one CSV record -> fields (RFC-4180 quotes, "" escape)
inline std::vector<std::string> split_csv_line(const std::string &line) {
    std::vector<std::string> out;
    std::string cur;
    bool quoted = false;
    for (size_t i = 0; i < line.size(); i++) {
        char ch = line[i];
        if (quoted) {
            if (ch == '"') {
                if (i + 1 < line.size() && line[i + 1] == '"') { cur.push_back('"'); i++; }
                else quoted = false;
            } else cur.push_back(ch);
        } else if (ch == '"') quoted = true;
        else if (ch == ',') { out.push_back(cur); cur.clear(); }
        else if (ch != '\r') cur.push_back(ch);
    }
    out.push_back(cur);
    return out;
}

inline CsvType infer_cell(CsvType so_far, const std::string &cell) {
    if (cell.empty()) return so_far;                  // null: no effect on the type
    long long l;
    double d;
    switch (so_far) {
    case CsvType::Null:
    case CsvType::Int:
        if (parse_long(cell, l)) {
            if (l >= INT32_MIN && l <= INT32_MAX) return so_far == CsvType::Null ? CsvType::Int : so_far;
            return CsvType::Long;
        }
        return parse_double(cell, d) ? CsvType::Double : CsvType::String;
    case CsvType::Long:
        if (parse_long(cell, l)) return CsvType::Long;
        return parse_double(cell, d) ? CsvType::Double : CsvType::String;
    case CsvType::Double:
        return parse_double(cell, d) ? CsvType::Double : CsvType::String;
    default:
        return CsvType::String;
    }
}

inline CsvTable read_csv(const std::string &path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::string line;
    if (!std::getline(f, line)) throw std::runtime_error("empty CSV (no header): " + path);
    CsvTable t;
    for (auto &h : split_csv_line(line)) {
        CsvColumn c;
        c.name = h;
        t.cols.push_back(std::move(c));
    }
    while (std::getline(f, line)) {
        if (line.empty() || line == "\r") continue;
        auto fields = split_csv_line(line);
        if (fields.size() != t.cols.size()) {
            std::ostringstream m;
            m << path << ": row " << (t.nrows + 2) << " has " << fields.size() << " fields, header has " << t.cols.size();
            throw std::runtime_error(m.str());
        }
        for (size_t c = 0; c < fields.size(); c++) t.cols[c].raw.push_back(std::move(fields[c]));
        t.nrows++;
    }
    // pass 1: infer (Spark runs a separate job for this); pass 2: convert
    for (auto &c : t.cols) {
        for (const auto &cell : c.raw) c.type = infer_cell(c.type, cell);
        if (c.type == CsvType::Null) c.type = CsvType::String;   // all-null column: Spark falls back to string
        c.validity.assign((size_t)((t.nrows + 7) / 8), 0);
        for (int64_t r = 0; r < t.nrows; r++) {
            const std::string &cell = c.raw[(size_t)r];
            const bool valid = !cell.empty();
            if (valid) c.validity[(size_t)(r >> 3)] |= (uint8_t)(1u << (r & 7));
            else c.has_nulls = true;
            long long l = 0;
            double d = 0.0;
            switch (c.type) {
            case CsvType::Int: if (valid) parse_long(cell, l); c.i32.push_back((int32_t)l); break;
            case CsvType::Long: if (valid) parse_long(cell, l); c.i64.push_back(l); break;
            case CsvType::Double: if (valid) parse_double(cell, d); c.f64.push_back(d); break;
            default: break;
            }
        }
        if (!c.has_nulls) c.validity.clear();
        if (c.type != CsvType::String) { c.raw.clear(); c.raw.shrink_to_fit(); }
    }
    return t;
}

} // namespace fsxhost
