// tour.hpp -- host-side mirror of the reference's feature-engineering job on top of the C ABI.
//
// Mirrors io.hops.examples.featurestore_tour.featuregroups.ComputeFeatures (ComputeFeatures.scala)
// and Main (Main.scala): same entry-point names, same argument meaning, same call order, same
// constants, same output schemas (case-class field order).  Where the Scala code calls Spark SQL
// (`groupBy/sum/count/join/map`) this code calls libfsx through include/fsx.h ONLY; where it calls
// hsfs (`fs.createFeatureGroup()...save(df)`) it hands the DataFrame and the statistics computed by
// libfsx to a FeatureGroupSink (hsfs persistence is out of scope; DirSink writes CSV + JSON).
// The reference's toolchain (Scala/JVM) is absent in this image, hence C++ (see INTEGRATION.md for
// the JNI binding a maintainer adds to keep the Scala signatures).
#pragma once
#include <cstdio>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/fsx.h"
#include "csv.hpp"

namespace fsxhost {

// ---- stand-ins for the reference's collaborators ------------------------------------------------
struct Logger {                        // org.apache.log4j.Logger
    bool quiet = false;
    void info(const std::string &m) const { if (!quiet) fprintf(stderr, "INFO %s\n", m.c_str()); }
};

struct HostColumn {
    std::string name;
    int32_t dtype = FSX_I32;
    int64_t n = 0;
    std::vector<unsigned char> data;
    std::vector<uint8_t> validity;     // empty = no nulls
    template <typename T> const T *as() const { return reinterpret_cast<const T *>(data.data()); }
    bool valid(int64_t i) const { return validity.empty() || ((validity[(size_t)(i >> 3)] >> (i & 7)) & 1); }
};

struct DataFrame {                     // what the Scala code hands to fg.save(...)
    std::vector<HostColumn> cols;
    int64_t nrows = 0;
    const HostColumn &col(const std::string &name) const;
};

struct FeatureGroupMeta {              // the builder chain of CF:108-116
    std::string name;
    int version = 1;
    std::string description;
    std::string time_travel_format = "NONE";
    std::vector<std::string> primary_keys, partition_keys;
    bool online_enabled = false;
    bool stats_enabled = true, stats_histograms = true, stats_correlations = true;   // StatisticsConfig(true,true,true)
};

struct FeatureGroupSink {              // fg.save(df) / td.save(query)
    virtual ~FeatureGroupSink() = default;
    virtual void save(const FeatureGroupMeta &meta, const DataFrame &df, const std::string &statistics_json) = 0;
    virtual const DataFrame &get(const std::string &name) const = 0;               // fs.getFeatureGroup(name)
};

class DirSink : public FeatureGroupSink {   // writes <dir>/<name>_<version>.csv and .statistics.json
  public:
    explicit DirSink(std::string dir) : dir_(std::move(dir)) {}
    void save(const FeatureGroupMeta &meta, const DataFrame &df, const std::string &statistics_json) override;
    const DataFrame &get(const std::string &name) const override;
  private:
    std::string dir_;
    std::map<std::string, DataFrame> saved_;
};

// ---- the job ---------------------------------------------------------------------------------------
class ComputeFeatures {
  public:
    // Constants of ComputeFeatures.scala:19-34
    static constexpr const char *TEAMS_DATASET_FILE = "teams.csv";
    static constexpr const char *TEAMS_FEATUREGROUP = "teams_features";
    static constexpr const char *GAMES_DATASET_FILE = "games.csv";
    static constexpr const char *GAMES_FEATUREGROUP = "games_features";
    static constexpr const char *SEASON_FEATUREGROUP_TOUR_ON_DEMAND = "season_features_on_demand";
    static constexpr const char *GAMES_FEATUREGROUP_TOUR_HUDI = "games_features_hudi_tour";
    static constexpr const char *PLAYERS_DATASET_FILE = "players.csv";
    static constexpr const char *PLAYERS_FEATUREGROUP = "players_features";
    static constexpr const char *ATTENDANCES_DATASET_FILE = "attendances.csv";
    static constexpr const char *ATTENDANCES_FEATUREGROUP = "attendances_features";
    static constexpr const char *SEASON_SCORES_DATASET_FILE = "season_scores.csv";
    static constexpr const char *SEASON_SCORES_FEATUREGROUP = "season_scores_features";
    static constexpr const char *TOUR_TRAINING_DATASET = "tour_training_dataset_test";
    static constexpr int FEATUREGROUP_VERSION = 1;
    static constexpr const char *SLASH_DELIMITER = "/";

    ComputeFeatures(fsx_ctx *ctx, FeatureGroupSink *sink) : ctx_(ctx), fs_(sink) {}

    void computeGamesFeatureGroup(const Logger &log, const std::string &datasetPath);         // CF:101
    void computeSeasonScoresFeatureGroup(const Logger &log, const std::string &datasetPath);  // CF:142
    void computeSeasonScoresOnDemandFeatureGroup(const Logger &log);                          // CF:179
    void computeAttendanceFeatureGroup(const Logger &log, const std::string &datasetPath);    // CF:200
    void computePlayersFeatureGroup(const Logger &log, const std::string &datasetPath);       // CF:239
    void computeTeamsFeatureGroup(const Logger &log, const std::string &datasetPath);         // CF:286
    void createTrainingDataset(const Logger &log);                                            // CF:312

    // Statistics of a set of device columns as the Deequ-style JSON hsfs posts
    // dictionaries: for a string column the device column holds dictionary codes (fsx_dictionary_encode_utf8, nulls as
    // validity) and dictionaries[name] the distinct strings: it is profiled as Deequ profiles a categorical column
    // (completeness, approximate distinct count, value -> count histogram), not as numbers.
    std::string computeStatistics(const std::vector<std::pair<std::string, const fsx_col *>> &cols,
                                  bool histograms, bool correlations,
                                  const std::map<std::string, std::vector<std::string>> &dictionaries = {});
    // every column of a headered CSV file (Spark-style inferred schema), strings included, as the same JSON
    std::string profileCsv(const std::string &path, bool histograms = true, bool correlations = true);
    // every column of a Parquet table (one file, or the part files of a directory in order) that libfsx decodes on the GPU
    // (fsx_parquet_decode_column_files / _strings_files: numeric columns, dictionary-encoded strings); columns it declines
    // (PLAIN-stored strings, ...) are listed in *skipped and left out
    std::string profileParquet(const std::vector<std::string> &paths, std::vector<std::string> *skipped = nullptr, bool histograms = true,
                               bool correlations = true);

  private:
    fsx_ctx *ctx_;
    FeatureGroupSink *fs_;
};

// spark.read.format("delta").load(dir): the part files the table's transaction log (_delta_log/*.json) leaves active, in
// the order they were added, as paths under `dir`
std::vector<std::string> deltaActiveFiles(const std::string &dir);

// Main.scala:60-66 -- expands a relative input with hdfs:///Projects/<project>/; absolute local paths
// and URIs pass through (this build reads the local file system only).
std::string parseInput(const std::string &input, const std::string &projectName);

// Main.scala:25-52: the seven calls in the reference's order.
void runTourJob(fsx_ctx *ctx, const std::string &input, FeatureGroupSink *sink, const Logger &log);

} // namespace fsxhost
