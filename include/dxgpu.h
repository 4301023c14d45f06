/*
 * dxgpu.h — C ABI of libdxgpu.so, the B200 (sm_100a) flow processor for Data Accelerator.
 *
 * This is the drop-in boundary: the functions a JNI / Java-FFM shim inside the reference's
 * Scala host (datax-host) binds in place of the Spark-SQL work done by
 * CommonProcessorFactory.createProcessor (reference:
 * DataProcessing/datax-host/src/main/scala/datax/processor/CommonProcessorFactory.scala:50-657).
 * Plain pointers and sizes only; no C++ or torch types cross this line.  INTEGRATION.md shows
 * the Scala-side binding.
 *
 * Conventions: every function returning int returns 0 on success, non-zero on failure (the
 * message is available through dx_ctx_last_error); nothing here falls back to a CPU path —
 * an unsupported construct is an error at create time.  Device pointers handed out by the
 * library stay valid until the owning object is released.
 */
#ifndef DXGPU_H
#define DXGPU_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct dx_ctx dx_ctx;         /* one per (process, GPU): stream, memory pool, error text   */
typedef struct dx_decoder dx_decoder; /* K1: JSON events -> columns for one input schema           */
typedef struct dx_table dx_table;     /* device-resident columnar table                            */

/* column physical types (dx_table_column_type) */
#define DX_TYPE_I64 1   /* long/int/short/byte                                  8 B/row */
#define DX_TYPE_F64 2   /* double/float                                         8 B/row */
#define DX_TYPE_STR 3   /* dictionary code (int32) + per-column dictionary      4 B/row */
#define DX_TYPE_BOOL 4  /* int64 0/1                                            8 B/row */
#define DX_TYPE_TS 5    /* timestamp, int64 microseconds since epoch            8 B/row */
#define DX_TYPE_FLAGS 6 /* rule flag bitmap, dx_table_flag_words uint32 per row         */

#define DX_MEM_HOST 0
#define DX_MEM_DEVICE 1

/* ---- context -------------------------------------------------------------------------- */
/* Replaces the SparkSession / executor pool a processor runs on
 * (reference: datax/host/SparkSessionSingleton.scala:20-24).                                */
int dx_ctx_create(int device_ordinal, dx_ctx** out);
void dx_ctx_destroy(dx_ctx* ctx);
const char* dx_ctx_last_error(const dx_ctx* ctx);
/* CUDA-event timing of every kernel stage on the context's stream (off by default). */
int dx_ctx_set_timing(dx_ctx* ctx, int enabled);
/* Sum of recorded milliseconds and number of samples for a stage ("k1_decode", ...); resets it. */
int dx_ctx_take_stage_ms(dx_ctx* ctx, const char* stage, double* total_ms, int64_t* samples);
/* Number of launches of this library's kernels since the last call (resets the counter). */
int64_t dx_ctx_take_launches(dx_ctx* ctx);
/* Work counters for roofline arithmetic ("k3_rows": rows visited by the aggregation kernel, "gather_bytes"); resets. */
int64_t dx_ctx_take_counter(dx_ctx* ctx, const char* name);
void* dx_ctx_stream(dx_ctx* ctx);                     /* cudaStream_t, for event timing by the caller */
int dx_ctx_synchronize(dx_ctx* ctx);
/* pooled device memory for callers that stage inputs themselves */
int dx_dev_alloc(dx_ctx* ctx, int64_t bytes, void** out);
void dx_dev_free(dx_ctx* ctx, void* p);
/* page-locked host memory, so that the copies inside process_batch / decode run at link speed; the pages are taken from the
 * NUMA node the context's GPU is attached to when the platform reports one (dx_ctx_numa_node: -1 if it does not) */
int dx_host_alloc(dx_ctx* ctx, int64_t bytes, void** out);
int dx_ctx_numa_node(dx_ctx* ctx);
void dx_host_free(dx_ctx* ctx, void* p);
int dx_copy_d2h_async(dx_ctx* ctx, void* dst_host, const void* src_dev, int64_t bytes);   /* on the ctx stream */
int dx_copy_h2d(dx_ctx* ctx, void* dst_dev, const void* src_host, int64_t bytes);
int dx_copy_d2h(dx_ctx* ctx, void* dst_host, const void* src_dev, int64_t bytes);

/* ---- K1: JSON decode -------------------------------------------------------------------- */
/* Replaces from_json(Raw, rawBlockSchema) in project()
 * (reference: CommonProcessorFactory.scala:93; schema JSON as read by
 * datax/input/SchemaFile.scala:22-34 — a Spark StructType in JSON).
 * Leaves of nested structs become columns named by their dotted path.                       */
int dx_decoder_create(dx_ctx* ctx, const char* spark_schema_json, dx_decoder** out);
void dx_decoder_destroy(dx_decoder* d);
/* json: n_bytes of events.  offsets == NULL: NDJSON framing (every '\n' ends a record, a final
 * unterminated record counts).  offsets != NULL: n_events+1 int64 offsets, record i =
 * json[offsets[i], offsets[i+1]).  n_events < 0 (NDJSON only): counted on the device.
 * memory_kind says where json/offsets live.  The resulting table owns its columns.           */
int dx_decoder_decode(dx_decoder* d, const uint8_t* json, int64_t n_bytes, const int64_t* offsets,
                      int64_t n_events, int memory_kind, dx_table** out);
/* No counterpart in the reference: an entry point for hosts that must not allocate device memory once
 * the job is running.  Typical cases: a Spark executor sharing the GPU with other flows, where a
 * cudaMalloc/cudaFree inside the per-batch path would synchronise the whole device and stall every
 * other stream (and could fail late under memory pressure); or a latency-bound ingest loop that
 * pre-sizes all of its buffers at start-up.  Same decode as dx_decoder_decode on NDJSON-framed device
 * input, but every output is caller-owned and the kernel's scratch (tile counter, statistics, per-tile
 * look-back states) lives in a fixed array inside the library image: about 512 KiB of device memory on
 * every GPU that loads libdxgpu.so, held until the process exits.  The call performs no cudaMalloc,
 * cudaFree or pooled allocation.
 *   json_dev, n_bytes : the batch in DEVICE memory; n_events must be stated (>= 0).
 *   column_ptrs[n_columns] : one device buffer per leaf column, in the order and with the element
 *       width of the columns of a table returned by dx_decoder_decode (8 B per row; 4 B for strings).
 *   mask_ptr : n_events validity elements, 1 byte each when the schema has <= 7 leaf columns, else
 *       4 bytes: bit k = column k non-NULL, top bit = the event parsed to a non-NULL struct.
 *   rows / null_rows : optional counts for this call.
 * String columns receive int32 dictionary codes; the text behind a code is read with
 * dx_decoder_dict_size / dx_decoder_dict_string below (codes are per decoder and only ever grow).
 * Limit: one call handles at most 65536 tiles of 32 KiB = 2 GiB of JSON; larger batches are rejected
 * with an error (split them, or use dx_decoder_decode).  The tile size is fixed at 32 KiB here because
 * the scratch capacity is stated in tiles; dx_decoder_decode picks the tile from the mean record length
 * and is a few percent faster on long records.
 * The scratch array is shared by all calls on a device and still holds the previous call's values, so
 * each call clears the part it uses first; calls whose GPU work overlaps (other streams, other
 * threads) race on it and must be serialised by the caller.  The first call on a device also looks up
 * the array's device address (a one-time runtime query): make that first call before any stream
 * capture is begun.  Every call waits for its kernel and reads the row counts back before it returns.
 * The calling thread's current CUDA device is the same after the call as before it.                */
int dx_decoder_decode_noalloc(dx_decoder* d, const uint8_t* json_dev, int64_t n_bytes, int64_t n_events,
                              void* const* column_ptrs, int n_columns, void* mask_ptr,
                              int64_t* rows, int64_t* null_rows);
/* The decoder's dictionary of a string leaf column (column index as in the decoded table): number of codes handed out
 * so far, and the UTF-8 text of one code.  For callers that hold only the code arrays dx_decoder_decode_noalloc wrote;
 * a dx_table answers the same questions through dx_table_dict_size / dx_table_dict_string.  These two copy the new
 * dictionary entries to the host on demand (no device allocation).                                                */
int dx_decoder_dict_size(dx_decoder* d, int column, int64_t* n_codes);
int dx_decoder_dict_string(dx_decoder* d, int column, int64_t code, const char** bytes, int64_t* len);
/* rows decoded / rows that became NULL structs / rows taken by the template fast path, last call */
int dx_decoder_stats(const dx_decoder* d, int64_t* rows, int64_t* null_rows, int64_t* fast_rows);
/* Records of the last call that Spark's from_json would have parsed but this decoder turned into NULL structs because they
 * exceed one of its per-record limits (an object or array under a string field, an escaped string longer than
 * 256 bytes, a number of more than 19 significant digits that sits exactly between two doubles).  They are counted in
 * null_rows as well; the flow reports them as the metric Input_Unparsed_Events_Count.  Such a record never fails the batch. */
int64_t dx_decoder_limited_rows(const dx_decoder* d);

/* ---- tables ----------------------------------------------------------------------------- */
/* What the Scala shim wraps into a DataFrame for the unchanged sinks
 * (reference: OutputOperator.output(DataFrame, Timestamp), datax-core sink/package.scala:32-35). */
int64_t dx_table_num_rows(const dx_table* t);
int dx_table_num_columns(const dx_table* t);
const char* dx_table_column_name(const dx_table* t, int col);
int dx_table_column_type(const dx_table* t, int col);
/* Leaves of a MAP('k', v, ...) / STRUCT(v AS n, ...) select item (the alert shaping CodegenRules emits:
 * DataX.Flow.CodegenRules.Tests/CGenMixedAlert.txt:45-68 "pivots"/"aggs"/"result", CGenNoCode.txt:75-99) are
 * stored as one column per leaf named <alias>.<key>[.<key>...].  Returns the number of levels below the alias
 * (0 for an ordinary column, -1 for a bad index); *kinds gets one 'm' or 's' per level, dx_table_column_path the
 * key at that level, so a sink can rebuild the nested value (map keys may themselves contain dots).          */
int dx_table_column_nesting(const dx_table* t, int col, const char** kinds);
const char* dx_table_column_path(const dx_table* t, int col, int level);
int dx_table_find_column(const dx_table* t, const char* name);      /* -1 if absent */
int dx_table_flag_words(const dx_table* t, int col);                /* FLAGS columns */
/* Raw device views: data pointer (NULL for constant columns), validity mask pointer with element
 * width (1 or 4 bytes) and the bit holding this column's validity (mask NULL => never NULL).  */
int dx_table_column_device(const dx_table* t, int col, const void** data, const void** vmask,
                           int* vmask_width, int* vmask_bit);
/* Copies a column to host: data = n_rows elements of the physical width (FLAGS: words*4 per row),
 * valid = n_rows bytes (1 = non-NULL).  Either destination may be NULL.                        */
int dx_table_column_to_host(const dx_table* t, int col, void* data, uint8_t* valid);
/* Dictionary of a STR column: number of codes, and the UTF-8 text of one code.                */
int dx_table_dict_size(const dx_table* t, int col, int64_t* n_codes);
int dx_table_dict_string(const dx_table* t, int col, int64_t code, const char** bytes, int64_t* len);
void dx_table_release(dx_table* t);

/* Rule metadata of a FLAGS column: the constant MAP('ruleId',...,'severity',...) of rule r as a JSON
 * object.  Flags are stored word-major: flags[w * n_rows + row], bit (r % 32) of word r / 32 set
 * <=> rule r's condition is TRUE for the row — the device form of the reference's
 * filterNull(Array(IF(cond_r, MAP(...), NULL), ...)) AS Rules (CodegenRules Engine.cs:232-266).   */
int dx_table_rule_count(const dx_table* t, int col);
int dx_table_rule_json(const dx_table* t, int col, int rule, const char** json, int64_t* len);
/* Rows for which each rule fired (SURVEY.md 8(d) C2 "flags + per-rule counts"; what a sink would get from
 * SELECT explode(Rules) ... GROUP BY ruleId): counts[r], r < dx_table_rule_count, computed on the device.       */
int dx_table_rule_hits(const dx_table* t, int col, int64_t* counts, int n_counts);

/* ---- design time: the rules / flow code generator ------------------------------------------ */
/* Engine.GenerateCode(code, rules, queryTemplates, outputTemplates, productId) of DataX.Flow.CodegenRules (Engine.cs:32-107): the
 * user's code (ProcessRules / ProcessAlerts / ProcessAggregateRules / ProcessAggregateAlerts / CreateMetric calls, OUTPUT ... TO,
 * CREATE TABLE, WITH UPSERT, TIMEWINDOW) + the rule definitions (JSON) + the two template files -> the transform text
 * (`--DataXQuery--` statements, RulesCode.Code) that dx_flow_create takes, byte for byte what the reference's generator emits
 * (its 18 test expectations are this repo's tests).  info_out (optional) receives one line per side result:
 * "output<TAB>table<TAB>sink", "table<TAB>name<TAB>schema" (CREATE TABLE), "window<TAB>view<TAB>duration", "metric<TAB>table".
 * On failure code_out holds the message.  No GPU is involved.                                                              */
int dx_codegen_rules(const char* user_code, const char* rules_json, const char* query_templates_xml, const char* output_templates_xml,
                     const char* product_id, char* code_out, int64_t code_cap, char* info_out, int64_t info_cap);
/* The same from a flow's GUI JSON, as the config generation does it (S450_PrepareTransformFile.cs:48-64: gui.process.queries joined by
 * newlines; RuleDefinitionGenerator.cs:15-33: gui.rules[].properties with $productId = the flow's name; property names stored as "_S_x"
 * are read as "$x").  A rule without a $ruleId gets "<flow name>-rule-<position>" (the reference draws a GUID).                 */
int dx_codegen_flow(const char* flow_json, const char* query_templates_xml, const char* output_templates_xml, char* code_out,
                    int64_t code_cap, char* info_out, int64_t info_cap);

/* ---- state tables on disk ------------------------------------------------------------------- */
/* StateTableHandler keeps every accumulator table as two Hive tables `STORED AS PARQUET LOCATION '<location>/A/'` and `.../B/'` plus
 * `<location>/metadata.info` holding `active=A|B` and `standby=B|A` (datax/handler/StateTableHandler.scala:26-52, 70-83); a batch
 * INSERT-OVERWRITEs the standby copy, flips and persists the metadata (:99-125).  dx_flow_state_save writes the table's ACTIVE copy
 * as one Parquet file into `<location>/<active>/` (the directory's other files are removed, as INSERT OVERWRITE does) and the
 * metadata file; dx_flow_state_load reads metadata.info and every part file of the active directory — files written by
 * dx_flow_state_save or by Spark 2.4 (v1 data pages, PLAIN or dictionary encoded, UNCOMPRESSED or SNAPPY, INT96 timestamps) — into
 * the flow's table; an absent location is an empty table with active=A, as CREATE TABLE IF NOT EXISTS gives.  Local paths only:
 * moving the files to and from HDFS / blob storage stays with the JVM side.                                             */
struct dx_flow;
int dx_flow_state_save(struct dx_flow* flow, const char* table, const char* location);
int dx_flow_state_load(struct dx_flow* flow, const char* table, const char* location);
/* The same file format on host columns (column_data: int64 / double bits / boolean 0|1 / timestamp microseconds per row; strings:
 * n_rows + 1 int64 offsets into string_bytes; column_valid: one byte per row or NULL = no NULLs).  No GPU is involved.      */
typedef struct dx_parquet dx_parquet;
int dx_parquet_write(const char* path, int n_columns, const char* const* names, const int* types, int64_t n_rows,
                     const void* const* column_data, const uint8_t* const* column_valid, const uint8_t* const* string_bytes,
                     char* err, int64_t err_cap);
int dx_parquet_read(const char* path_or_dir, dx_parquet** out, char* err, int64_t err_cap);
int64_t dx_parquet_num_rows(const dx_parquet* p);
int dx_parquet_num_columns(const dx_parquet* p);
const char* dx_parquet_column_name(const dx_parquet* p, int column);
int dx_parquet_column_type(const dx_parquet* p, int column);
int dx_parquet_column(const dx_parquet* p, int column, const void** data, const uint8_t** valid, const uint8_t** string_bytes);
void dx_parquet_release(dx_parquet* p);

/* ---- blob input: gzip on the device ------------------------------------------------------- */
/* The blob input's files are gzip-compressed JSON lines (datax/input/BlobPointerInput.scala; the lines reach project() through
 * CPF:463-515).  dx_gunzip_device inflates the n_members gzip members that lie back to back in gz (member m = bytes
 * [member_offsets[m], member_offsets[m+1]); one member per file, RFC 1952; host memory for memory_kind 0, device for 1) into ONE
 * device buffer holding their contents in order — ready for dx_flow_process_batch(..., memory_kind = device).  Output offsets
 * come from the members' ISIZE trailers and the decoded length is checked against them; the trailer's CRC-32 is not verified.
 * *out_dev is pool memory: release it with dx_dev_free.                                                                  */
int dx_gunzip_device(dx_ctx* ctx, const uint8_t* gz, const int64_t* member_offsets, int n_members, int memory_kind,
                     void** out_dev, int64_t* n_bytes);

/* ---- sink hand-off: rows as JSON ---------------------------------------------------------- */
/* What OutputManager.sink builds before it calls the JSON sinks (datax-host sink/OutputManager.scala:103-118):
 * to_json(struct(<every column whose name does not start with "__DataX_">)) for each row of an output table, produced on the
 * device as NDJSON: row r is bytes[offsets[r] .. offsets[r+1] - 1) and is followed by one '\n' (so the byte string is also what
 * BlobSinker writes, sink/BlobSinker.scala:54-92, before gzip).  Serialisation rules (Spark 2.4.5 JacksonGenerator restated):
 * NULL struct fields are skipped, map values are written as null; doubles as java.lang.Double.toString prints them (shortest
 * round-trip digits; NaN and the infinities quoted); strings with Jackson's default escapes; timestamps as
 * yyyy-MM-dd'T'HH:mm:ss.SSSXXX in UTC; the values of one MAP(...) in Spark's common type; a rule-flag column as the array of the
 * fired rules' objects (filterNull(Array(IF(cond, MAP(...), NULL), ...)), CodegenRules Engine.cs:232-266).  The handle owns the
 * device buffers until dx_json_release.                                                                                     */
typedef struct dx_json dx_json;
int dx_table_to_json(const dx_table* t, dx_json** out);
int64_t dx_json_num_rows(const dx_json* j);
int64_t dx_json_num_bytes(const dx_json* j);
int dx_json_device(const dx_json* j, const uint8_t** bytes, const int64_t** offsets);     /* device pointers; offsets has n_rows + 1 entries */
int dx_json_to_host(const dx_json* j, uint8_t* bytes, int64_t* offsets);                  /* either destination may be NULL */
void dx_json_release(dx_json* j);
/* One sink's filter column (OutputManager.scala:53-61, 110-113: `<filter expr> AS _<prefix>_<sink>`, e.g. conf.filter of the
 * EventHub / HTTP / SQL sinks, blobConf.groupEvaluation): flags[r] = 1 where the SQL boolean expression is TRUE for row r, 0 where
 * it is FALSE or NULL (SinkerUtil.outputFilteredEvents keeps a row when the flag is not NULL and true, SinkerUtil.scala:31).   */
int dx_table_filter_flags(const dx_table* t, const char* expr, uint8_t* flags);

/* ---- the flow processor (drop-in boundary) ------------------------------------------------ */
typedef struct dx_flow dx_flow;       /* one per flow: decoder, parsed SQL, window cache, state tables */
typedef struct dx_result dx_result;   /* what one micro-batch produced                                */
/* Replaces CommonProcessorFactory.createProcessor(config) (CommonProcessorFactory.scala:50-126).
 * conf_text: contents of <flow>.conf (ConfigManager.scala:98-135); schema_json: inputschema.json;
 * projection_steps: contents of each projection file (ProjectionHandler.scala:23-36);
 * transform_text: <flow>-combined.txt (TransformHandler.scala:25-53).  Unsupported SQL, types or
 * hooks fail here, at construction, with a message — never later and never on a CPU path.       */
int dx_flow_create(dx_ctx* ctx, const char* conf_text, const char* schema_json,
                   const char* const* projection_steps, int n_steps, const char* transform_text,
                   dx_flow** out);
void dx_flow_destroy(dx_flow* flow);
/* Replaces CommonProcessor.processDataset(data, batchTime, batchInterval, outputPartitionTime,
 * targets, namespace) (CommonProcessorFactory.scala:333-399): decode -> projection -> windows ->
 * SQL statements -> state tables.  now_ms is the fixed clock behind current_timestamp().
 * Calls on one (flow, name_space) must not overlap.                                               */
int dx_flow_process_batch(dx_flow* flow, const char* name_space, const uint8_t* json, int64_t n_bytes,
                          const int64_t* offsets, int64_t n_events, int memory_kind,
                          int64_t batch_time_ms, int64_t interval_ms, int64_t output_partition_time_ms,
                          int64_t now_ms, dx_result** out);
/* The frame the reference's ingest adapters build around every event (CommonProcessorFactory.scala:405-421, 458-461, 628-654):
 * beside Raw, the column __DataX_FileInfo (datax/data/FileInternal.scala:14-20: inputPath, outputFolders, outputFileName,
 * fileTime, ruleIndexPrefix, target), kept through every projection step (CPF:95-100), and, visible to the projection
 * expressions, DataXProperties (PropertiesHandler.scala:26-35: BatchTime, BlobTime, BlobPathHint, CPTime, CPExecutor plus the
 * conf's appendproperty.* entries) and the transport's Properties / SystemProperties maps.  Here they are per-BATCH constants:
 * without a call every field of __DataX_FileInfo is NULL (what the streaming adapters pass, FileInternal()) and the two
 * transport maps are empty; a blob or EventHub shim calls dx_flow_set_batch_info before dx_flow_process_batch with the values
 * of the partition it is about to hand over (per-event property values are not carried).  Result tables show the struct as one
 * column per leaf (dx_table_column_nesting: __DataX_FileInfo.inputPath ..., outputFolders.<group>).  info == NULL resets.      */
typedef struct dx_batch_info {
    const char* input_path;                       /* NULL = SQL NULL, as for every string below */
    const char* const* output_folder_keys; const char* const* output_folder_values; int n_output_folders;   /* < 0: NULL map */
    const char* output_file_name;
    int64_t file_time_ms; int has_file_time;
    const char* rule_index_prefix;
    const char* target;
    const char* const* property_keys; const char* const* property_values; int n_properties;
    const char* const* system_property_keys; const char* const* system_property_values; int n_system_properties;
    const char* executor_id;                      /* DataXProperties.CPExecutor; NULL = "driver" */
} dx_batch_info;
int dx_flow_set_batch_info(dx_flow* flow, const dx_batch_info* info);
/* Per-EVENT transport properties: the Properties (which = 0) or SystemProperties (which = 1) value of one key for every event of the
 * NEXT batch, as the EventHub and Kafka adapters put them beside each body (processEventData CPF:405-422: Row(body, properties,
 * systemProperties, ...); processConsumerRecord CPF:628-655).  bytes + offsets (n_events + 1, ascending) hold the UTF-8 values;
 * valid[i] == 0 (valid may be NULL: all present) means event i does not carry the key (SQL NULL).  The values become a string
 * column `Properties.<key>` / `SystemProperties.<key>` for that batch (replacing a per-batch constant of the same key from
 * dx_flow_set_batch_info) and are consumed by the next process call, whose event count must equal n_events.            */
int dx_flow_set_event_property(dx_flow* flow, int which, const char* key, const uint8_t* bytes, const int64_t* offsets,
                               const uint8_t* valid, int64_t n_events);

/* Ingest that runs ahead of the processor.  In the reference the receivers fill the next micro-batch while the job thread
 * works on the current one (Spark Streaming's receiver / job split behind StreamingHost.scala:62-95); over PCIe that overlap
 * is worth having, because copying a batch of JSON takes far longer than processing it.  dx_flow_stage_batch starts the
 * host->device copy of one NDJSON batch on a side stream and returns at once; dx_flow_process_staged is
 * dx_flow_process_batch on that batch (it makes the processor's stream wait for the copy on the device, consumes the
 * handle and returns its device buffer to the pool).  json_host must stay unchanged until dx_flow_process_staged returns
 * and should be page-locked (dx_host_alloc), or the copy is neither asynchronous nor at link speed.  A handle that is
 * never processed is dropped with dx_staged_release.                                                                  */
typedef struct dx_staged dx_staged;
int dx_flow_stage_batch(dx_flow* flow, const uint8_t* json_host, int64_t n_bytes, dx_staged** out);
int dx_flow_process_staged(dx_flow* flow, const char* name_space, dx_staged* staged, int64_t n_events,
                           int64_t batch_time_ms, int64_t interval_ms, int64_t output_partition_time_ms,
                           int64_t now_ms, dx_result** out);
void dx_staged_release(dx_staged* staged);
/* The Map[String, Double] processDataset returns (Input_Normalized_Events_Count, ...). */
int dx_result_metrics(dx_result* r, const char* const** keys, const double** vals, int* n);
int dx_result_num_tables(const dx_result* r);
const char* dx_result_table_name(const dx_result* r, int i);
/* A named dataset of this batch (OutputManager looks tables up by name, CPF:297-306); the returned
 * handle must be released with dx_table_release and may outlive the result.                       */
int dx_result_table(dx_result* r, const char* dataset, dx_table** out);
void dx_result_release(dx_result* r);
/* Accumulator (state) tables: the active physical table and whether it is the "B" copy
 * (StateTableHandler.scala:26-52: metadata.info active=/standby=), for Parquet A/B persistence. */
int dx_flow_state_export(dx_flow* flow, const char* table, dx_table** out, int* active_is_b);
/* Restart: what the JVM read back from the active Parquet copy (StateTableHandler.scala:70-96 registers the active
 * table at start; SURVEY.md 8(b) dx_flow_state_import) becomes the active physical table again.  Host buffers, one per
 * column in the order of the table's declared schema (datax.job.process.statetable.<name>.schema): 8-byte values for
 * long / double / timestamp (microseconds) / boolean columns; for string columns int64 offsets[n_rows + 1] into
 * string_bytes[k].  column_valid[k]: one byte per row (0 = NULL), or NULL for a column without NULLs.           */
int dx_flow_state_import(dx_flow* flow, const char* table, int64_t n_rows, int n_columns, const void* const* column_data,
                         const uint8_t* const* column_valid, const uint8_t* const* string_bytes, int active_is_b);

/* Reference data (ReferenceDataHandler.scala:42-60, CSVUtil.scala:27-29): the CSV text of
 * datax.job.input.default.referencedata.<name>.path becomes a view <name> whose columns are all
 * strings; joins against numeric event columns cast the string side, as Spark does.  Call after
 * dx_flow_create and before the first batch.                                                      */
int dx_flow_add_refdata(dx_flow* flow, const char* name, const char* csv_text, char delimiter, int has_header);
/* User functions (datax-core extension/DynamicUDF.scala:12-46, SparkJarLoader.scala:82-155): a
 * String -> String function callable from the flow SQL.  User code cannot run on the GPU: the library
 * calls fn once per DISTINCT argument value (dictionary entry) and keeps the results as a device lookup
 * table.  Return 0 on success; *out_len = -1 yields NULL; the result is copied out of `out` (out_cap bytes). */
typedef int (*dx_udf1_fn)(void* user, const char* arg, int64_t arg_len, char* out, int64_t out_cap, int64_t* out_len);
int dx_flow_register_udf(dx_flow* flow, const char* name, dx_udf1_fn fn, void* user);
/* The other arities (DynamicUDF.Generator0..3, Java UDF0..UDF3 by class name, SparkJarLoader.scala:82-155; Azure-function
 * UDFs of 0-3 arguments, AzureFunctionHandler.scala:42-64): a function of n_args in 0..3 values returning a string.
 * Arguments reach the callback as text — strings as they are, integers in decimal, timestamps as Spark prints them — and the
 * library calls it once per DISTINCT combination of argument values (once per batch for n_args == 0), keeping the answers
 * as a device lookup table.  A NULL argument makes the result NULL without a call when n_args >= 2.  refresh_each_batch != 0
 * is for functions whose answers change with time (a DynamicUDF with an onInterval handler): nothing is kept from one batch
 * to the next.  Doubles cannot be passed (no decimal rendering of doubles here): CAST them in the SQL or fail at first use.  */
typedef int (*dx_udfn_fn)(void* user, int n_args, const char* const* args, const int64_t* arg_lens, char* out, int64_t out_cap, int64_t* out_len);
int dx_flow_register_udfn(dx_flow* flow, const char* name, int n_args, dx_udfn_fn fn, void* user, int refresh_each_batch);
/* IntervalUpdateHandler.onInterval (DynamicUDF.scala:40-46, called for every dynamic UDF at the start of each batch,
 * CommonProcessorFactory.scala:351-353): the host refreshes whatever its functions depend on.                              */
typedef void (*dx_interval_fn)(void* user, int64_t batch_time_ms);
int dx_flow_set_interval_hook(dx_flow* flow, dx_interval_fn fn, void* user);
/* StringNormalizer (datax-core extension/StringNormalizer.scala:10-12; applied to Raw before from_json, CPF:82, 93; sample:
 * datax-udf-samples RemoveInvalidChars).  User code runs on the host: before decoding a batch that arrives in HOST memory the
 * library hands the whole NDJSON batch to fn, which writes the normalised batch (the same records, one per line) to out
 * (out_cap = 2 * n_in + 64 bytes) and returns 0.  The time it takes is reported as the metric Latency-Normalizer.  A batch in
 * device memory cannot be normalised: with a normalizer set such a call fails.  fn == NULL removes it.                     */
typedef int (*dx_normalizer_fn)(void* user, const uint8_t* in, int64_t n_in, uint8_t* out, int64_t out_cap, int64_t* n_out);
int dx_flow_set_normalizer(dx_flow* flow, dx_normalizer_fn fn, void* user);

/* ---- one 8xB200 box: event-range sharding + key-hash shuffle ------------------------------- */
/* Spark runs every GROUP BY as partial aggregate -> Exchange(hashpartitioning(keys)) -> final aggregate
 * (the plans behind spark.sql at CommonProcessorFactory.scala:271).  Here each GPU decodes, projects
 * and rule-checks the events it received; the rows that feed keyed window aggregates are sent to the
 * GPU that owns hash64(group key) % world, whose flow keeps that key's window cache and state.
 * The exchange itself is one NCCL all-to-all issued by the host (one process per GPU).           */
/* project() only (CPF:90-103): decode + projection steps, no window/SQL work. */
int dx_flow_project(dx_flow* flow, const uint8_t* json, int64_t n_bytes, const int64_t* offsets,
                    int64_t n_events, int memory_kind, int64_t now_ms, dx_table** out);
/* route() only (CPF:131-328) on an already projected table, e.g. the rows received from the shuffle. */
int dx_flow_process_table(dx_flow* flow, const char* name_space, const dx_table* projected,
                          int64_t batch_time_ms, int64_t interval_ms, int64_t now_ms, dx_result** out);
/* K5 pack: writes the named columns of every row as fixed records (8 B per column + 8 B validity bits,
 * *rec_bytes) into packed_dev (device, n_rows * rec_bytes, caller-owned), grouped by owner rank;
 * counts[r] = rows for rank r.  key_cols must be among cols.  Dictionary-coded columns are rank-local
 * and are refused.                                                                               */
int dx_table_shuffle_pack(const dx_table* t, const char* const* cols, int n_cols, const char* const* key_cols,
                          int n_keys, int world, void* packed_dev, int64_t* counts, int64_t* rec_bytes);
/* K5 fused with the exchange (one kernel does pack + transfer over peer memory instead of pack, NCCL all-to-all):
 * dx_table_shuffle_count gives the rows per owner; the ranks exchange the counts (a few integers), each derives where
 * its block starts in every owner's receive buffer, and dx_table_shuffle_put writes every record straight into the
 * owner's buffer through peer pointers (peer_bufs[r] = rank r's buffer as mapped here, first_record[r] = first record
 * index of this rank's block there).  The receive buffers are ordinary dx_dev_alloc memory made visible to the other
 * processes with dx_ipc_export / dx_ipc_open (CUDA IPC, NVLink peer access); a barrier between the puts and the
 * owners' dx_table_from_packed is the caller's.  Same Exchange(hashpartitioning) as above.                       */
int dx_table_shuffle_count(const dx_table* t, const char* const* cols, int n_cols, const char* const* key_cols,
                           int n_keys, int world, int64_t* counts);
int dx_table_shuffle_put(const dx_table* t, const char* const* cols, int n_cols, const char* const* key_cols,
                         int n_keys, int world, void* const* peer_bufs, const int64_t* first_record);
/* The same put without the count exchange before it: every sender owns a fixed block in every owner's receive buffer
 * (first_record[r] = where this rank's block starts in rank r's buffer, block_records[r] = how many records it holds), so
 * the ranks exchange their per-owner counts (sent[r]) once AFTER the puts — the one collective that also tells an owner that
 * every put has landed — and the owner unpacks the blocks with dx_table_from_packed_blocks.  A block that turns out too small
 * is an error of the call (nothing is written past it).                                                                  */
int dx_table_shuffle_put_blocks(const dx_table* t, const char* const* cols, int n_cols, const char* const* key_cols,
                                int n_keys, int world, void* const* peer_bufs, const int64_t* first_record,
                                const int64_t* block_records, int64_t* sent);
int dx_ipc_export(dx_ctx* ctx, const void* dev_ptr, uint8_t* handle64);      /* 64-byte handle for another process */
int dx_ipc_open(dx_ctx* ctx, const uint8_t* handle64, void** dev_ptr);        /* map a peer's buffer */
int dx_ipc_close(dx_ctx* ctx, void* dev_ptr);
/* K5 unpack: n_rows received records -> a table with the given column names and DX_TYPE_* types. */
int dx_table_from_packed(dx_ctx* ctx, const char* const* col_names, const int* col_types, int n_cols,
                         const void* packed_dev, int64_t n_rows, dx_table** out);
/* ... from n_blocks runs of the buffer (block_first / block_count in records, one block per sender), in block order. */
int dx_table_from_packed_blocks(dx_ctx* ctx, const char* const* col_names, const int* col_types, int n_cols,
                                const void* packed_dev, int n_blocks, const int64_t* block_first,
                                const int64_t* block_count, dx_table** out);

/* ---- host-side artefact parsers, exposed for tests (no GPU needed) ------------------------- */
/* Canonical text of a statement as the SQL-subset parser understood it; error text on failure. */
int dx_sql_canonical(const char* sql, char* buf, int64_t cap);
/* TransformSQLParser.parse (TransformSqlParser.scala:27-95): one line per command
 * "name<TAB>type<TAB>text", then "#ref<TAB>name<TAB>count" lines.                               */
int dx_transform_split(const char* text, char* buf, int64_t cap);
/* stringToTimestamp (ConcurrentDateFormat.scala:38-62), UTC session: 1 and *us on success, 0 = NULL. */
int dx_string_to_timestamp(const char* s, int64_t* us);
/* scala Duration.create(...).toSeconds as used by TimeWindowHandler.scala:33-35; -1 on error.     */
int64_t dx_duration_seconds(const char* s);

/* ---- synthetic events (measurement tooling; shapes of the reference's simulated IoT data) -- */
/* Event shapes after DataGenerator.scala:27-209 / HomeAutomationLocal.json:288. */
typedef struct dx_gen_params {
    uint64_t seed;
    int64_t n_keys;          /* deviceId uniform in 1..n_keys                    */
    int64_t n_homes;         /* homeId = 1 + mix(deviceId) % n_homes             */
    int64_t base_ms;         /* batch time, epoch milliseconds                   */
    int64_t interval_ms;
    int64_t watermark_ms;
    int32_t shape;           /* 1 = S1 (5 fields), 2 = S2 (+temperature,humidity) */
    int32_t late_permille;   /* events older than the watermark                  */
    int32_t future_permille; /* events stamped after the batch time              */
    int32_t pad;
} dx_gen_params;
/* Writes events [first, first+n) as NDJSON into a pooled device buffer (*bytes, *n_bytes);
 * offsets (optional) receives a device array of n+1 offsets.  Free both with dx_dev_free.      */
int dx_gen_events(dx_ctx* ctx, const dx_gen_params* p, int64_t first, int64_t n, uint8_t** bytes,
                  int64_t* n_bytes, int64_t** offsets);

#ifdef __cplusplus
}
#endif
#endif /* DXGPU_H */
