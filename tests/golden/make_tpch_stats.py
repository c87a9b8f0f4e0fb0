"""Extracts the lineitem statistics the reference ships for its tpch connector (run in the build container, where
/root/reference exists) into tests/golden/tpch_lineitem_stats.json."""
import json
import os

SRC = "/root/reference/plugin/trino-tpch/src/main/resources/tpch/statistics"
COLUMNS = ["l_orderkey", "l_partkey", "l_linenumber", "l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_returnflag", "l_linestatus",
           "l_shipdate", "l_commitdate", "l_receiptdate"]
out = {}
for sf in ("sf0.01", "sf1.0"):
    d = json.load(open(os.path.join(SRC, sf, "lineitem.json")))
    out[sf] = {"rowCount": d["rowCount"],
               "columns": {c: {k: d["columns"][c][k] for k in ("distinctValuesCount", "min", "max")} for c in COLUMNS}}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tpch_lineitem_stats.json"), "w"), indent=1, sort_keys=True)
