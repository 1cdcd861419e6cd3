"""CPU: the host layer's wire decode / encode (host/klm_wire.h) against the string-level
oracle: the same items and specs the Go controller would see after decoding the API
server's JSON, including apimachinery's int64-for-integral-numbers quirk."""
import json

import controller_oracle as co
import helpers

KATS = helpers.load_kats()


def test_decode_uav_items_matches_oracle(pkg):
    s = pkg.hostbinding.HostSession(None)
    items = [c for case in KATS for c in case["items"]]
    items.append({"spec": {"node_name": 7, "battery": {"remaining_percent": "90"}}, "status": {"collection_status": 3}})
    items.append({"spec": {"node_name": 'esc"é\n', "battery": {"remaining_percent": 1e2}}})
    items.append({})
    assert s.load_uavs_json(json.dumps({"items": items}, ensure_ascii=False)) == len(items)
    bat, flags, names = helpers.flatten_items(items)
    for i, it in enumerate(items):
        u = s.uav_info(i)
        assert u["battery"] == bat[i]
        spec = it.get("spec", {})
        want_name = spec.get("node_name") if isinstance(spec.get("node_name"), str) else ""
        assert u["name"] == want_name and u["has_name"] == isinstance(spec.get("node_name"), str)
        assert u["status"].lower() == co.read_string(it.get("status", {}), "collection_status").lower()
    assert s.load_uavs_json(json.dumps(items[:3])) == 3          # a bare array is accepted too
    s.close()


def test_decode_requests_int_quirk(pkg):
    s = pkg.hostbinding.HostSession(None)
    specs = [c["spec"] for c in KATS]
    specs.append({"workload": {"name": "w", "namespace": "d"}, "minBatteryPercent": 80, "preferredNodes": ["a", 5, "B"]})
    specs.append({"workload": {"name": "w", "namespace": "d"}, "minBatteryPercent": 8e1})
    items = [{"spec": sp, "status": {"phase": ph}} if ph is not None else {"spec": sp}
             for sp, ph in zip(specs, [None, "Pending", "Assigned"] * len(specs))]
    text = json.dumps({"items": items}, ensure_ascii=False)
    # json.dumps writes 50.0 for Python floats and 80 for ints, like the YAML examples' integers
    for quirk in (True, False):
        assert s.load_requests_json(text, apimachinery_ints=quirk) == len(items)
        for i, (sp, it) in enumerate(zip(specs, items)):
            want = co.decode_spec(sp, apimachinery_ints=quirk)
            got = s.request_info(i)
            assert got["min_battery"] == want.min_battery_percent, (i, quirk)
            assert got["n_pref"] == len(want.preferred_nodes)
            assert got["workload"] == want.workload.name
            assert got["has_phase"] == ("status" in it)
    s.close()


def test_bad_json_is_an_error(pkg):
    s = pkg.hostbinding.HostSession(None)
    for bad in ("", "{", '{"items": [1,]}', '{"items": {}}', '[{"spec": {"node_name": "x"}}] tail'):
        try:
            s.load_uavs_json(bad)
        except RuntimeError as e:
            assert "JSON" in str(e)
        else:
            raise AssertionError(f"accepted {bad!r}")
    s.close()


def test_json_float_matches_go_encoding(pkg):
    """`score` in the status map is written like Go's encoding/json does (controller.go:233 puts a
    float64 into the unstructured map): shortest round-trip digits, plain decimal for exponents in
    [-6, 21), exponent form outside with e-0X trimmed to e-X. Expected strings are Go's output."""
    go = {90.0: "90", 100.0: "100", 110.0: "110", 85.5: "85.5", 0.0: "0", 0.1: "0.1", 95.2: "95.2",
          1e20: "100000000000000000000", 1e21: "1e+21", 1e-6: "0.000001", 1e-7: "1e-7", 1.5e-10: "1.5e-10",
          123456789.125: "123456789.125", -3.25: "-3.25", 1.2345678901234567e19: "12345678901234567000",
          5e-324: "5e-324", 1.7976931348623157e308: "1.7976931348623157e+308", 100.10000000000001: "100.10000000000001"}
    for v, want in go.items():
        assert pkg.hostbinding.format_json_float(v) == want, v
    # and it always parses back to the same double
    import random
    rnd = random.Random(5)
    for _ in range(2000):
        v = rnd.choice([rnd.uniform(0, 110), rnd.uniform(-1e6, 1e6), rnd.random() * 10 ** rnd.randint(-30, 30),
                        round(rnd.uniform(0, 110), 1)])
        assert float(pkg.hostbinding.format_json_float(v)) == v
