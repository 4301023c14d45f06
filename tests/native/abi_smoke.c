/* abi_smoke.c — a plain C99 program against include/dxgpu.h and libdxgpu.so, the way a foreign-function binding sees them:
 * only C types cross the boundary.  It calls the entry points that need no GPU (the artefact parsers) and prints their answers;
 * tests/test_abi.py builds and runs it. */
#include <stdio.h>
#include <string.h>
#include "dxgpu.h"

int main(void) {
    static char buf[1 << 16];
    int64_t us = 0;
    if (dx_duration_seconds("5 minutes") != 300) { puts("duration"); return 1; }
    if (dx_string_to_timestamp("2018-08-10T22:55:03Z", &us) != 1 || us != 1533941703000000LL) { printf("timestamp %lld\n", (long long)us); return 1; }
    if (dx_sql_canonical("select a, count(*) as n from t where b > 1 group by a", buf, (int64_t)sizeof buf) != 0) { printf("sql: %s\n", buf); return 1; }
    printf("%s\n", buf);
    if (dx_transform_split("--DataXQuery--\nT1 = SELECT * FROM x\n", buf, (int64_t)sizeof buf) != 0) { printf("split: %s\n", buf); return 1; }
    printf("%s", buf);
    if (dx_sql_canonical("select from", buf, (int64_t)sizeof buf) == 0) { puts("bad SQL accepted"); return 1; }
    printf("error text: %s\n", buf);
    return 0;
}
