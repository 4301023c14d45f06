// dx_flow.h — the per-micro-batch flow processor behind the C ABI (dx_flow_*).
//
// Host-side restatement target: the closure built by CommonProcessorFactory.createProcessor
// (reference: datax/processor/CommonProcessorFactory.scala:50-657): project() :90-103,
// route() :131-328, processDataset() :333-399, with TimeWindowHandler.scala:41-67,
// StateTableHandler.scala:70-128, TransformSqlParser.scala:27-104, ConfigManager.scala:98-135.
#pragma once
#include "dx_common.h"
#include "dx_sql.h"
#include "dx_expr.h"
#include "dx_agg.h"
#include "dx_aggp.h"
#include "dx_decode_host.h"
#include "dx_strings.h"
#include <set>
#include <chrono>

struct SqlCommand {
    std::string text, name;        // name empty => "Command" type (TransformSqlParser.scala:24-26)
    bool is_query = false;
    dxsql::Statement stmt;
};
struct ParsedTransform {
    std::vector<SqlCommand> commands;
    std::map<std::string, int> view_reference_count;
};
// TransformSQLParser.parse restated (reference: datax/sql/TransformSqlParser.scala:27-95)
int dx_parse_transform(const std::string& text, ParsedTransform* out, std::string* err);

struct RuleTable {
    std::vector<std::vector<std::pair<std::string, std::string>>> rules;   // MAP('k','v',...) per rule
    bool single = false;   // a lone IF(cond, MAP(...), NULL): the column is a nullable map, not filterNull(Array(...))'s array
};

struct CiLess {   // Spark resolves table (and column) names case-insensitively (spark.sql.caseSensitive=false)
    bool operator()(const std::string& a, const std::string& b) const {
        size_t n = a.size() < b.size() ? a.size() : b.size();
        for (size_t i = 0; i < n; i++) { int x = tolower((unsigned char)a[i]), y = tolower((unsigned char)b[i]); if (x != y) return x < y; }
        return a.size() < b.size();
    }
};
struct View {
    std::vector<DxTableP> chunks;
    bool ranged = false;
    std::string ts_col;
    long long lo_us = 0, hi_us = 0;
    long long pane_us = 0;          // batch interval: windows whose ends are multiples of it can be served from per-batch pane partials
};

// Per cached batch and statement: the batch's rows pre-aggregated by (group keys, pane = floor(ts / interval)).
// A window aggregate then combines one partial per cached batch instead of rescanning every cached event each interval
// (the reference rescans: CommonProcessorFactory.scala:203-227 re-filters the union of all cached batches per window).
struct PanePartialP { DxTableP chunk; AggPPartial ap; };
struct PanePartial {
    DxTableP chunk;                 // keeps the batch (and with it the cache key, its address) alive
    std::string sig;                // keys / values / accumulators / pane width this partial was built for
    std::vector<DevBufP> keys;      // n_keys + 1 arrays of int64 (last = pane)
    DevBufP valid32;                // bit k set: key k not NULL
    std::vector<DevBufP> accs;      // raw accumulators, one array per accumulator
    int64_t n = 0;
};

typedef std::map<std::string, View, CiLess> ViewMap;

struct StateTable {
    std::string name;
    std::vector<std::pair<std::string, int>> schema;
    DxTableP phys[2];              // "A" and "B"
    int active = 0;                // metadata.info active=A standby=B initially (StateTableHandler.scala:45-50)
    bool flipped = false;
};

struct NamespaceState {
    std::map<long long, DxTableP> past_batches;    // pastRdds: arrival batch time (ms) -> cached rows
};

struct dx_result {
    struct dx_flow* flow = nullptr;
    std::map<std::string, DxTableP> datasets;
    std::vector<std::string> dataset_order;
    std::vector<std::string> metric_keys;
    std::vector<double> metric_vals;
    std::vector<const char*> metric_key_ptrs;
};

bool dx_is_metadata_column(const std::string& name);      // __DataX_*, DataXProperties.*, Properties.*, SystemProperties.*

struct dx_flow {
    dx_ctx* ctx = nullptr;
    std::map<std::string, std::string> conf;
    std::unique_ptr<dx_decoder> decoder;
    std::vector<std::vector<dxsql::SelectItem>> projections;     // one list of selectExpr items per step
    ParsedTransform transform;
    // time windows (TimeWindowHandler.scala:15-21)
    std::vector<std::pair<std::string, long long>> windows;       // name -> seconds
    bool windows_enabled = false;
    std::string ts_column;
    long long watermark_s = 0, max_window_s = 0;
    std::map<std::string, StateTable, CiLess> state_tables;
    std::map<std::string, DxTableP, CiLess> ref_tables;           // reference data views (ReferenceDataHandler.scala:42-60)
    std::vector<std::string> outputs;
    // transport metadata of the batch being processed (dx_flow_set_batch_info): the FileInternal struct behind __DataX_FileInfo
    // (datax/data/FileInternal.scala:14-20; all NULL for the streaming adapters, CPF:417, 459) and the per-batch maps
    // per-EVENT values of one Properties / SystemProperties key for the next batch (dx_flow_set_event_property): the maps the
    // EventHub and Kafka adapters build per record (CPF:405-421, 628-654)
    struct EventProperty { std::string column; std::vector<int32_t> codes; std::vector<uint32_t> valid; };
    std::vector<EventProperty> event_props;                   // consumed by the next project()
    std::map<std::string, DxDictP> event_prop_dicts;          // one dictionary per column for the life of the flow (codes stay comparable across batches)
    int set_event_property(int which, const char* key, const uint8_t* bytes, const int64_t* offsets, const uint8_t* valid, int64_t n_events);
    int add_event_property_columns(DxTable* t);
    struct BatchInfo {
        bool has_input_path = false, has_output_folders = false, has_output_file_name = false, has_file_time = false,
             has_rule_index_prefix = false, has_target = false;
        std::string input_path, output_file_name, rule_index_prefix, target, executor_id = "driver";
        long long file_time_ms = 0;
        std::vector<std::pair<std::string, std::string>> output_folders, properties, system_properties;
    } batch_info;
    std::vector<std::pair<std::string, std::string>> append_properties;   // conf datax.job.process.appendproperty.* (PropertiesHandler.scala:22-24)
    std::map<std::string, NamespaceState> ns;
    std::map<std::string, unsigned long long> agg_capacity_hint;  // per statement name
    std::map<std::string, std::map<const DxTable*, PanePartial>> pane_partials;   // partial signature -> cached batch -> partial
    std::map<std::string, std::map<const DxTable*, PanePartialP>> pane_partials_p;   // same, partitioned shared-memory form (dx_aggp.h)
    std::map<std::string, int> pane_P;       // partitions per signature (every partial of a signature shares it)
    // the partial shapes built so far for one (window source, pane width, group keys) family; a statement uses any member
    // whose accumulators cover its own
    struct PaneVariant { std::vector<dxsql::ExprP> vals; std::vector<AggAccDesc> accs; std::vector<std::string> desc; std::string sig; };
    std::map<std::string, std::vector<PaneVariant>> pane_variants;
    int pane_variant_seq = 0;
    std::map<std::string, std::pair<long long, long long>> pane_fill;   // per signature: (fullest partition, rows) of the last build
    int pane_partitions_first = 0;  // conf datax.job.process.gpu.panepartitions: first partition count tried (tests force the re-partition path)
    bool pane_smem_off = false;     // conf datax.job.process.gpu.panesmem=false: partials through the global hash table (dx_agg.cu)
    std::map<std::string, std::shared_ptr<void>> rule_plans;        // (statement, item) -> planned rule set with its device tables (dx_rules.h)
    bool pane_partials_off = false; // conf datax.job.process.gpu.panepartials=false: always rescan (tests compare both)
    std::string last_error;
    // user functions evaluated on the host side of the boundary, once per distinct dictionary entry
    typedef int (*udf1_fn)(void* user, const char* arg, int64_t arg_len, char* out, int64_t out_cap, int64_t* out_len);
    typedef int (*udfn_fn)(void* user, int n_args, const char* const* args, const int64_t* arg_lens, char* out, int64_t out_cap, int64_t* out_len);
    // n_args 0..3; fn (one string argument) or fnn (any arity); refresh: results are not kept from batch to batch (DynamicUDF
    // with an onInterval handler, DynamicUDF.scala:12-46: the function may answer differently after every refresh)
    struct Udf { udf1_fn fn = nullptr; void* user = nullptr; int n_args = 1; udfn_fn fnn = nullptr; bool refresh = false; };
    typedef void (*interval_fn)(void* user, int64_t batch_time_ms);
    interval_fn on_interval = nullptr; void* on_interval_user = nullptr;          // CPF:351-353: udfs.foreach(onInterval(spark, batchTime))
    typedef int (*normalizer_fn)(void* user, const uint8_t* in, int64_t n_in, uint8_t* out, int64_t out_cap, int64_t* n_out);
    normalizer_fn normalizer = nullptr; void* normalizer_user = nullptr;          // StringNormalizer before from_json (CPF:82, 93)
    std::vector<uint8_t> normalized;                                              // the normalised batch of the call in progress
    double normalizer_seconds = 0;
    std::map<std::string, Udf> udfs;                               // upper-cased name -> callback
    std::map<std::string, std::shared_ptr<LutCache>> luts;        // (expression, dictionary) -> device lookup table
    std::map<std::string, DxDictP> expr_dicts;                     // (statement, item) -> dictionary of a string-valued result, stable across batches
    std::chrono::steady_clock::time_point batch_t0; bool batch_t0_set = false;   // start of the batch being processed (Latency-Process)
    long long pending_unparsed = 0;    // decoder->last_stats.limited_rows of the batch being routed
    int rules_mode = 0;                // DX_RULES_* bits from conf datax.job.process.gpu.rulesstaging / .rulesfast (dx_rules.h)
    bool generic_rules_only = false;   // tests: force the bytecode interpreter for rule arrays (conf datax.job.process.gpu.genericrules=true)
    int project(const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events, int memory_kind, long long now_ms, DxTableP* out, long long batch_ms = -1);
    void add_metadata_columns(DxTable* t, long long batch_ms, long long now_ms) const;
    int route(const std::string& name_space, const DxTableP& projected, long long batch_ms, long long interval_ms, long long now_ms, dx_result** out);
    int process(const std::string& name_space, const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events,
                int memory_kind, long long batch_ms, long long interval_ms, long long now_ms, dx_result** out);
};

int dx_flow_build(dx_ctx* ctx, const char* conf_text, const char* schema_json, const char* const* projection_steps, int n_steps,
                  const char* transform_text, dx_flow** out);
// stringToTimestamp restated for dictionary entries (reference: ConcurrentDateFormat.scala:38-62)
bool dx_string_to_timestamp_us(const std::string& s, long long* out_us);
long long dx_parse_duration_seconds(const std::string& s, bool* ok);
int dx_flow_add_refdata_impl(dx_flow* F, const std::string& name, const std::string& csv, char delim, bool header);
