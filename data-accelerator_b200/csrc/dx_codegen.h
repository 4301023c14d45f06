// dx_codegen.h — rules / flow code generator restated (see dx_codegen.cu).
#pragma once
#include <string>
#include <utility>
#include <vector>
struct dx_codegen_result {
    std::string code;                                              // RulesCode.Code
    std::vector<std::pair<std::string, std::string>> outputs;      // (table, sink): RulesCode.Outputs
    std::vector<std::pair<std::string, std::string>> tables;       // (name, schema): RulesCode.AccumlationTables
    std::vector<std::pair<std::string, std::string>> windows;      // (view, duration): RulesCode.TimeWindows
    std::vector<std::string> metrics;                              // tables sent to the Metrics sink (MetricsRoot's sources)
};
int dx_codegen_impl(const std::string& user_code, const std::string& rules_json, const std::string& query_xml, const std::string& output_xml,
                    const std::string& product, dx_codegen_result* res, std::string* err);
int dx_codegen_flow_impl(const std::string& flow_json, const std::string& query_xml, const std::string& output_xml, dx_codegen_result* res, std::string* err);
