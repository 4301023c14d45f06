"""Adversarial JSON records for decode parity (missing fields, nulls, type
mismatches, escapes, shuffled key order, extra fields, whitespace, truncation)."""
import random

GOOD = '{"deviceDetails":{"deviceId":%d,"deviceType":"%s","eventTime":%d,"homeId":%d,"status":%d,"temperature":%s,"humidity":%s}}'

HAND = [
    '',
    '   ',
    '{}',
    '{"deviceDetails":null}',
    '{"deviceDetails":{}}',
    '{"deviceDetails":{"deviceId":1}}',
    '{"deviceDetails":{"deviceId":null,"deviceType":null}}',
    '{"deviceDetails":{"deviceId":"","status":1}}',
    '{"deviceDetails":{"deviceId":"12","status":1}}',
    '{"deviceDetails":{"deviceId":1.0,"status":1}}',
    '{"deviceDetails":{"deviceId":1e3,"status":1}}',
    '{"deviceDetails":{"deviceId":9223372036854775807,"status":-9223372036854775808}}',
    '{"deviceDetails":{"deviceId":9223372036854775808}}',
    '{"deviceDetails":{"deviceId":-9223372036854775809}}',
    '{"deviceDetails":{"deviceId":123456789012345678901234567890}}',
    '{"deviceDetails":{"deviceId":007}}',
    '{"deviceDetails":{"deviceId":-0,"temperature":-0}}',
    '{"deviceDetails":{"deviceId":+1}}',
    '{"deviceDetails":{"deviceId":true}}',
    '{"deviceDetails":{"deviceId":[1]}}',
    '{"deviceDetails":{"deviceId":{"a":1}}}',
    '{"deviceDetails":{"temperature":1,"humidity":-2}}',
    '{"deviceDetails":{"temperature":"NaN","humidity":"Infinity"}}',
    '{"deviceDetails":{"temperature":"-Infinity","humidity":"nan"}}',
    '{"deviceDetails":{"temperature":NaN,"humidity":-Infinity}}',
    '{"deviceDetails":{"temperature":Infinity,"humidity":INF}}',
    '{"deviceDetails":{"temperature":-INF}}',
    '{"deviceDetails":{"temperature":""}}',
    '{"deviceDetails":{"temperature":"1.5"}}',
    '{"deviceDetails":{"temperature":1.7976931348623157e308,"humidity":4.9e-324}}',
    '{"deviceDetails":{"temperature":1e309,"humidity":1e-400}}',
    '{"deviceDetails":{"temperature":0.1e1,"humidity":123456789012345678.5}}',
    '{"deviceDetails":{"temperature":2.2250738585072011e-308,"humidity":9007199254740993}}',
    '{"deviceDetails":{"temperature":0.30000000000000004,"humidity":1.0000000000000002}}',
    '{"deviceDetails":{"temperature":1.,"humidity":1}}',
    '{"deviceDetails":{"temperature":.5}}',
    '{"deviceDetails":{"temperature":1e,"humidity":1}}',
    '{"deviceDetails":{"temperature":-.5}}',
    '{"deviceDetails":{"deviceType":123}}',
    '{"deviceDetails":{"deviceType":1.50}}',
    '{"deviceDetails":{"deviceType":true}}',
    '{"deviceDetails":{"deviceType":""}}',
    '{"deviceDetails":{"deviceType":"a\\"b\\\\c\\/d\\b\\f\\n\\r\\t"}}',
    '{"deviceDetails":{"deviceType":"\\u00e9\\u20ac\\ud83d\\ude00"}}',
    '{"deviceDetails":{"deviceType":"\\ud83d"}}',
    '{"deviceDetails":{"deviceType":"\\x41"}}',
    '{"deviceDetails":{"deviceType":"tab\there"}}',
    '{"deviceDetails":{"deviceType":"é€ plain utf8"}}',
    "{'deviceDetails':{'deviceType':'single quoted','deviceId':5}}",
    '{"deviceDetails":{"deviceType":"a string longer than sixteen bytes for the hashed dictionary path"}}',
    '{"deviceDetails":{"deviceType":"escaped and longer than sixteen bytes \\n for the unescape buffer"}}',
    '{"deviceDetails":{"deviceId":1,"deviceId":2,"status":3,"status":null}}',
    '{"deviceDetails":{"deviceId":1},"deviceDetails":{"status":2}}',
    '{"deviceDetails":{"deviceId":1},"deviceDetails":null}',
    '{"DeviceDetails":{"deviceId":1}}',
    '{"deviceDetails":{"DeviceId":1,"deviceid":2}}',
    '{"extra":{"a":[1,2,{"b":null}],"c":"x"},"deviceDetails":{"deviceId":4,"more":[[],{}],"status":0}}',
    '{"extra":[1,2,,3],"deviceDetails":{"deviceId":4}}',
    '{"extra":{"a":tru},"deviceDetails":{"deviceId":4}}',
    '{"extra":"unterminated,"deviceDetails":{"deviceId":4}}',
    ' { "deviceDetails" : { "deviceId" : 7 , "deviceType" : "Heating" , "status" : 1 } } ',
    '{\t"deviceDetails":\r{"deviceId":8}}',
    '{"deviceDetails":{"deviceId":9}} trailing garbage',
    '{"deviceDetails":{"deviceId":9}}{"deviceDetails":{"deviceId":10}}',
    '{"deviceDetails":{"deviceId":9},}',
    '{"deviceDetails":{"deviceId":9,}}',
    '{"deviceDetails":{"deviceId":9}',
    '{"deviceDetails":{"deviceId":9',
    '{"deviceDetails":{"deviceId"',
    '{"deviceDetails":{"deviceId":}}',
    '{"deviceDetails":{"deviceId" 9}}',
    '{"deviceDetails":{deviceId:9}}',
    '{"deviceDetails":[{"deviceId":9}]}',
    '{"deviceDetails":"str"}',
    '{"deviceDetails":5}',
    '[{"deviceDetails":{"deviceId":11}}]',
    '[{"deviceDetails":{"deviceId":11}},{"deviceDetails":{"deviceId":12}}]',
    '[]',
    '[1]',
    '"just a string"',
    '123',
    'null',
    'true',
    '{"deviceDetails":{"status":1,"homeId":150,"eventTime":1600000000000,"deviceType":"Heating","deviceId":3}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":20.5,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":20.5,"humidity":40.25}} ',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":20,"humidity":40}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":2.05e1,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":20.1234567890123456,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Hea\\u0074ing","eventTime":1600000000000,"homeId":150,"status":1,"temperature":20.5,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":-3,"deviceType":"Heating","eventTime":-1,"homeId":0,"status":1,"temperature":-20.5,"humidity":-0.0}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":null,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":3,"deviceType":null,"eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":1234567890123456789,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":01,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":1,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":01.5,"humidity":40.25}}',
    '{"deviceDetails":{"deviceId":1,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":40.25}}}',
    '{"deviceDetails":{"deviceId":1,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":40.25}',
    '{"deviceDetails":{"deviceId":1,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":40.25,"x":1}}',
    '{"deviceDetails":{"deviceId":1,"deviceType":"Heating","eventTime":1600000000000,"homeId":150,"status":1,"temperature":1.5,"humidity":}}',
]


def random_number(rng):
    k = rng.randrange(8)
    if k == 0:
        return str(rng.randrange(-10**6, 10**6))
    if k == 1:
        return "%d.%0*d" % (rng.randrange(-1000, 1000), rng.randrange(1, 8), rng.randrange(10**6))
    if k == 2:
        return repr(rng.uniform(-1e6, 1e6))
    if k == 3:
        return "%se%d" % (repr(rng.uniform(1, 10)), rng.randrange(-320, 309))
    if k == 4:
        return "".join(rng.choice("0123456789") for _ in range(rng.randrange(1, 25))).lstrip("0") or "0"
    if k == 5:
        d = "".join(rng.choice("0123456789") for _ in range(rng.randrange(15, 30)))
        return "0." + d + "e%d" % rng.randrange(-300, 300)
    if k == 6:
        return repr(rng.random() * 10.0 ** rng.randrange(-300, 300))
    return "%d.%02d" % (rng.randrange(0, 100), rng.randrange(100))


def random_records(n, seed):
    """Mostly well-formed events with random perturbations."""
    rng = random.Random(seed)
    types = ["DoorLock", "WindowLock", "Heating", "GarageDoorLock", "Sp ace", "quo\\\"te", "ünï"]
    out = []
    for _ in range(n):
        rec = GOOD % (rng.randrange(1, 10**6), rng.choice(types), 1600000000000 + rng.randrange(10**9),
                      rng.randrange(1, 10**5), rng.randrange(2), random_number(rng), random_number(rng))
        k = rng.randrange(20)
        if k == 0:
            pos = rng.randrange(len(rec))
            rec = rec[:pos]                       # truncation
        elif k == 1:
            pos = rng.randrange(len(rec))
            rec = rec[:pos] + rng.choice('{}[]:,"x 0') + rec[pos + 1:]   # corruption
        elif k == 2:
            rec = rec.replace(",", " ,\t", 2).replace(":", " : ", 3)
        elif k == 3:
            rec = rec.replace('"status":', '"status":null,"ignored":')
        elif k == 4:
            rec = rec.replace('"homeId":', '"extra":{"deep":[1,{"x":"y"}]},"homeId":')
        out.append(rec)
    return out


def all_records(n_random=3000, seed=7):
    return HAND + random_records(n_random, seed)


def to_ndjson(records):
    """NDJSON framing cannot carry raw newlines inside a record; none of ours has one."""
    return ("\n".join(records) + "\n").encode("utf-8")
