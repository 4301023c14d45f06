"""Named cases for the Spark 2.4.5 semantics the oracle restates without a reference test behind them.
One entry per rule of medium or lower confidence, so that a later run against a real Spark 2.4.5 can confirm or correct each
by name (NEEDS_SPARK_CONFIRMATION.md lists them).  Every expectation is the WRITTEN rule, not an output of either
implementation; tests/test_spark_semantics.py checks the oracle against it, tests/test_spark_semantics_gpu.py the CUDA path."""
import math

NAN, INF = float("nan"), float("inf")

# name -> (record, row parsed to a non-NULL struct?, {leaf: expected value | None})        schema: S2 (dxo.iot_schema_json(2))
DECODE = {
    "long_from_float_token_nulls_row": ('{"deviceDetails":{"deviceId":1.0,"status":1}}', False, {"status": None}),
    "long_from_exponent_token_nulls_row": ('{"deviceDetails":{"deviceId":1e3,"status":1}}', False, {"status": None}),
    "long_from_nonempty_string_nulls_row": ('{"deviceDetails":{"deviceId":"12","status":1}}', False, {"status": None}),
    "long_from_empty_string_nulls_field_only": ('{"deviceDetails":{"deviceId":"","status":1}}', True, {"deviceId": None, "status": 1}),
    "long_from_boolean_token_nulls_row": ('{"deviceDetails":{"deviceId":true,"status":1}}', False, {"status": None}),
    "long_overflow_nulls_row": ('{"deviceDetails":{"deviceId":9223372036854775808,"status":1}}', False, {"status": None}),
    "long_min_value_parses": ('{"deviceDetails":{"deviceId":-9223372036854775808}}', True, {"deviceId": -2 ** 63}),
    "long_leading_zero_rejected": ('{"deviceDetails":{"deviceId":007}}', False, {}),
    "long_plus_sign_rejected": ('{"deviceDetails":{"deviceId":+1}}', False, {}),
    "double_from_integer_token": ('{"deviceDetails":{"temperature":5,"humidity":-2}}', True, {"temperature": 5.0, "humidity": -2.0}),
    "double_from_exact_case_nan_and_infinity_strings": ('{"deviceDetails":{"temperature":"NaN","humidity":"-Infinity"}}', True, {"temperature": NAN, "humidity": -INF}),
    "double_from_lowercase_nan_string_nulls_row": ('{"deviceDetails":{"temperature":"nan","status":1}}', False, {"status": None}),
    "double_from_numeric_string_nulls_row": ('{"deviceDetails":{"temperature":"1.5","status":1}}', False, {"status": None}),
    "double_from_empty_string_nulls_row": ('{"deviceDetails":{"temperature":"","status":1}}', False, {"status": None}),
    "double_bare_nan_and_inf_tokens_accepted": ('{"deviceDetails":{"temperature":NaN,"humidity":-INF}}', True, {"temperature": NAN, "humidity": -INF}),
    "double_overflow_and_underflow_round_to_inf_and_zero": ('{"deviceDetails":{"temperature":1e309,"humidity":1e-400}}', True, {"temperature": INF, "humidity": 0.0}),
    "double_integer_token_minus_zero_is_plus_zero": ('{"deviceDetails":{"temperature":-0}}', True, {"temperature": 0.0}),
    "double_float_token_minus_zero_keeps_sign": ('{"deviceDetails":{"temperature":-0.0}}', True, {"temperature": -0.0}),
    "string_from_number_token_keeps_its_text": ('{"deviceDetails":{"deviceType":1.50}}', True, {"deviceType": b"1.50"}),
    "string_from_boolean_token_keeps_its_text": ('{"deviceDetails":{"deviceType":true}}', True, {"deviceType": b"true"}),
    "string_empty_stays_empty_not_null": ('{"deviceDetails":{"deviceType":""}}', True, {"deviceType": b""}),
    "string_lone_surrogate_becomes_question_mark": ('{"deviceDetails":{"deviceType":"\\ud83d"}}', True, {"deviceType": b"?"}),
    "string_backslash_any_rejected": ('{"deviceDetails":{"deviceType":"\\x41"}}', False, {}),
    "string_unescaped_control_character_rejected": ('{"deviceDetails":{"deviceType":"tab\there"}}', False, {}),
    "single_quoted_names_and_strings_allowed": ("{'deviceDetails':{'deviceType':'q','deviceId':5}}", True, {"deviceType": b"q", "deviceId": 5}),
    "duplicate_key_last_wins_even_null": ('{"deviceDetails":{"deviceId":1,"deviceId":2,"status":3,"status":null}}', True, {"deviceId": 2, "status": None}),
    "field_names_are_case_sensitive": ('{"DeviceDetails":{"deviceId":1}}', True, {"deviceId": None}),
    "unknown_fields_are_skipped_but_validated": ('{"extra":[1,2,,3],"deviceDetails":{"deviceId":4}}', False, {}),
    "bytes_after_the_first_value_ignored": ('{"deviceDetails":{"deviceId":9}} trailing garbage', True, {"deviceId": 9}),
    "trailing_comma_rejected": ('{"deviceDetails":{"deviceId":9,}}', False, {}),
    "top_level_array_of_one_object_is_that_row": ('[{"deviceDetails":{"deviceId":11}}]', True, {"deviceId": 11}),
    "top_level_array_of_two_objects_is_null": ('[{"deviceDetails":{"deviceId":11}},{"deviceDetails":{"deviceId":12}}]', False, {}),
    "empty_and_blank_records_are_null_rows": ('   ', False, {}),
}


def ev(**kw):
    import json
    d = {"deviceId": 1, "deviceType": "DoorLock", "eventTime": 1_700_000_039_000, "homeId": 1, "status": 1, "temperature": 20.5, "humidity": 40.25}
    d.update(kw)
    d = {k: v for k, v in d.items() if v is not ...}
    body = json.dumps({"deviceDetails": d}, separators=(",", ":"))
    return body.replace('"@NaN"', "NaN")


# name -> (events, transform SQL (one statement named T), expected rows as a sorted list of tuples)
FLOW = {
    "comparison_with_null_is_null_and_where_keeps_true_only": (
        [ev(deviceId=1, status=None), ev(deviceId=2, status=0), ev(deviceId=3, status=1)],
        "T = SELECT deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.status <> 1 OR deviceDetails.status = 1",
        [(2,), (3,)]),
    "kleene_or_true_wins_over_null": (
        [ev(deviceId=1, status=None, homeId=7), ev(deviceId=2, status=None, homeId=8)],
        "T = SELECT deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.status = 1 OR deviceDetails.homeId = 7",
        [(1,)]),
    "nan_equals_nan_and_is_greater_than_every_number": (
        [ev(deviceId=1, temperature="@NaN"), ev(deviceId=2, temperature=1e300)],
        "T = SELECT deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.temperature > 1.0E305 AND deviceDetails.temperature = deviceDetails.temperature",
        [(1,)]),
    "long_column_against_decimal_literal_compares_as_double": (
        [ev(deviceId=1, homeId=10), ev(deviceId=2, homeId=11)],
        "T = SELECT deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.homeId > 10.5",
        [(2,)]),
    "division_is_double_and_by_zero_is_null": (
        [ev(deviceId=1, homeId=7, status=2), ev(deviceId=2, homeId=7, status=0)],
        "T = SELECT deviceDetails.deviceId, deviceDetails.homeId / deviceDetails.status AS q FROM DataXProcessedInput",
        [(1, 3.5), (2, None)]),
    "count_variants_and_aggregates_ignore_nulls": (
        [ev(deviceId=1, status=None, temperature=None), ev(deviceId=1, status=4, temperature=2.5), ev(deviceId=1, status=6, temperature=None)],
        "T = SELECT deviceDetails.deviceId, COUNT(*) AS n, COUNT(deviceDetails.status) AS ns, SUM(deviceDetails.status) AS s, MIN(deviceDetails.status) AS mn, "
        "MAX(deviceDetails.temperature) AS tx, AVG(deviceDetails.status) AS av FROM DataXProcessedInput GROUP BY deviceDetails.deviceId",
        [(1, 3, 2, 10, 4, 2.5, 5.0)]),
    "aggregates_over_only_nulls_are_null_but_counts_are_zero": (
        [ev(deviceId=1, status=None), ev(deviceId=1, status=None)],
        "T = SELECT deviceDetails.deviceId, COUNT(deviceDetails.status) AS ns, SUM(deviceDetails.status) AS s, MIN(deviceDetails.status) AS mn, AVG(deviceDetails.status) AS av "
        "FROM DataXProcessedInput GROUP BY deviceDetails.deviceId",
        [(1, 0, None, None, None)]),
    "sum_of_longs_wraps_around": (
        [ev(deviceId=1, homeId=9223372036854775807), ev(deviceId=1, homeId=1)],
        "T = SELECT deviceDetails.deviceId, SUM(deviceDetails.homeId) AS s FROM DataXProcessedInput GROUP BY deviceDetails.deviceId",
        [(1, -9223372036854775808)]),
    "min_max_put_nan_above_every_number": (
        [ev(deviceId=1, temperature="@NaN"), ev(deviceId=1, temperature=3.0), ev(deviceId=1, temperature=-1.0)],
        "T = SELECT deviceDetails.deviceId, MIN(deviceDetails.temperature) AS mn, MAX(deviceDetails.temperature) AS mx FROM DataXProcessedInput GROUP BY deviceDetails.deviceId",
        [(1, -1.0, NAN)]),
    "null_group_keys_form_one_group": (
        [ev(deviceId=None, status=1), ev(deviceId=None, status=2), ev(deviceId=5, status=3)],
        "T = SELECT deviceDetails.deviceId, SUM(deviceDetails.status) AS s FROM DataXProcessedInput GROUP BY deviceDetails.deviceId",
        [(None, 3), (5, 3)]),
    "minus_zero_and_zero_are_different_group_keys_in_2_4": (
        [ev(deviceId=1, temperature=-0.0), ev(deviceId=2, temperature=0.0), ev(deviceId=3, temperature=0.0)],
        "T = SELECT deviceDetails.temperature, COUNT(*) AS n FROM DataXProcessedInput GROUP BY deviceDetails.temperature",
        [(-0.0, 1), (0.0, 2)]),
    "global_aggregate_over_zero_rows_is_one_row": (
        [ev(deviceId=1)],
        "T = SELECT COUNT(*) AS n, SUM(deviceDetails.status) AS s FROM DataXProcessedInput WHERE deviceDetails.deviceId = 99",
        [(0, None)]),
    "grouped_aggregate_over_zero_rows_is_no_row": (
        [ev(deviceId=1)],
        "T = SELECT deviceDetails.deviceId, COUNT(*) AS n FROM DataXProcessedInput WHERE deviceDetails.deviceId = 99 GROUP BY deviceDetails.deviceId",
        []),
    "distinct_treats_nulls_as_equal": (
        [ev(deviceId=None, homeId=1), ev(deviceId=None, homeId=1), ev(deviceId=2, homeId=1)],
        "T = SELECT DISTINCT deviceDetails.deviceId, deviceDetails.homeId FROM DataXProcessedInput",
        [(None, 1), (2, 1)]),
    "string_equality_is_exact_bytes": (
        [ev(deviceId=1, deviceType="Lock"), ev(deviceId=2, deviceType="lock"), ev(deviceId=3, deviceType="Lock ")],
        "T = SELECT deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.deviceType = 'Lock'",
        [(1,)]),
    "like_percent_underscore_and_whole_string_match": (
        [ev(deviceId=1, deviceType="DoorLock"), ev(deviceId=2, deviceType="Door"), ev(deviceId=3, deviceType="xDoorLock"), ev(deviceId=4, deviceType=None)],
        "T = SELECT deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.deviceType LIKE 'Door_ock' OR deviceDetails.deviceType LIKE '%oor'",
        [(1,), (2,)]),
    "concat_is_null_if_any_argument_is_null_and_casts_longs": (
        [ev(deviceId=1, homeId=7), ev(deviceId=2, homeId=None)],
        "T = SELECT deviceDetails.deviceId, CONCAT('h', CAST(deviceDetails.homeId AS String)) AS c FROM DataXProcessedInput",
        [(1, b"h7"), (2, None)]),
    "hour_and_date_trunc_use_the_utc_session": (
        [ev(deviceId=1, eventTime=1_700_000_039_123)],
        "T = SELECT hour(CAST(deviceDetails.eventTime / 1000 AS timestamp)) AS h, DATE_TRUNC('second', CAST(deviceDetails.eventTime / 1000 AS timestamp)) AS t FROM DataXProcessedInput",
        [(22, 1_700_000_039_000_000)]),
    "current_timestamp_is_the_injected_clock": (
        [ev(deviceId=1)],
        "T = SELECT deviceDetails.deviceId, unix_timestamp() AS u FROM DataXProcessedInput",
        [(1, 1_700_000_041)]),
}


def same(a, b):
    if a is None or b is None:
        return a is None and b is None
    if isinstance(a, float) or isinstance(b, float):
        a, b = float(a), float(b)
        if math.isnan(a) or math.isnan(b):
            return math.isnan(a) and math.isnan(b)
        return a == b and math.copysign(1, a) == math.copysign(1, b)
    return a == b


def sort_key(row):
    return [(0, 0) if x is None else ((1, repr(x)) if isinstance(x, (bytes, str)) else (2, (0 if x != x else x), math.copysign(1, x) if isinstance(x, float) else 0)) for x in row]
