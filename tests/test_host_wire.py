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
