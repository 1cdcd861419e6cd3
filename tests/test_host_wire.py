"""CPU: the host layer's wire decode / encode (host/klm_wire.h) against the string-level
oracle: the same items and specs the Go controller would see after decoding the API
server's JSON, including apimachinery's int64-for-integral-numbers quirk."""
import json

import numpy as np

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


def _state_from_channels(c):
    """A UAVState JSON object (pkg/uav/mavlink_simulator.go:11-106) whose flattening is `c`."""
    modes = ["MANUAL", "STABILIZE", "LOITER", "AUTO", "RTL", "LAND"]
    missions = ["IDLE", "ACTIVE", "PAUSED", "COMPLETED"]
    health = ["OK", "WARNING", "CRITICAL", "ERROR"]
    f = lambda i: float(c[i])
    return {
        "uav_id": "uav-x", "node_name": "n", "system_time": "2026-01-01T00:00:00Z",
        "gps": {"latitude": f(0), "longitude": f(1), "altitude": f(2), "relative_altitude": f(3), "hdop": f(4),
                "satellite_count": int(c[5]), "fix_type": int(c[6]), "ground_speed": f(7),
                "course_over_ground": f(8), "timestamp": "2026-01-01T00:00:00Z"},
        "attitude": {"roll": f(9), "pitch": f(10), "yaw": f(11), "roll_rate": f(12), "pitch_rate": f(13),
                     "yaw_rate": f(14)},
        "flight": {"mode": modes[int(c[40])], "armed": bool(c[33]), "airspeed": f(15), "ground_speed": f(16),
                   "vertical_speed": f(17), "throttle_percent": f(18)},
        "battery": {"voltage": f(19), "current": f(20), "remaining_percent": f(21), "remaining_capacity": f(22),
                    "total_capacity": f(23), "temperature": f(24), "cell_count": int(c[25]),
                    "time_remaining": int(c[26])},
        "mission": {"current_waypoint": int(c[27]), "total_waypoints": int(c[28]),
                    "mission_state": missions[int(c[41])], "distance_to_wp": f(29), "eta_to_wp": int(c[30])},
        "health": {"system_status": health[int(c[42])],
                   "sensors_health": {k: bool(c[34 + i]) for i, k in enumerate(
                       ["accelerometer", "barometer", "battery", "compass", "gps", "gyroscope"])},
                   "error_count": int(c[31]), "warning_count": int(c[32]), "messages": []},
    }


def test_uav_report_decode_flatten_and_defaults(pkg):
    """models.UAVReport (models.go:182-192) -> the 64 channels in the order of SURVEY.md 8a A11,
    which is the order synth.uav_channel_rows generates; defaulting of manager.go:392-409."""
    s = pkg.hostbinding.HostSession(None)
    rows = pkg.synth.uav_channel_rows(3, 0, 3, 40).reshape(3, 64, 40)
    for node in range(3):
        for t in (0, 17, 39):
            c = rows[node, :, t]
            rep = {"node_name": f"node-{node}", "uav_id": f"uav-{node}", "source": "", "status": "",
                   "timestamp": "0001-01-01T00:00:00Z", "state": _state_from_channels(c)}
            d = s.report_decode(json.dumps(rep))
            assert d["has_state"] and d["node_name"] == f"node-{node}"
            assert d["status"] == "active" and d["source"] == "agent"           # manager.go:401-409
            assert np.array_equal(d["channels"].view(np.uint32), c.view(np.uint32)), (node, t)
            assert np.float32(d["remaining_percent"]) == c[21]                     # the scheduler's battery
    d = s.report_decode(json.dumps({"node_name": "", "uav_id": "u"}))              # dropped, manager.go:392-394
    assert d["node_name"] == "" and not d["has_state"]
    d = s.report_decode(json.dumps({"node_name": "n", "status": "degraded", "source": "pull",
                                    "state": {"flight": {"mode": "ACRO"}, "health": {"system_status": "??"}}}))
    assert d["status"] == "degraded" and d["source"] == "pull" and d["has_state"]
    assert d["channels"][40] == -1 and d["channels"][42] == -1 and d["channels"][41] == -1
    for bad in ('{"node_name": 5}', '{"node_name": "n", "state": {"gps": {"satellite_count": 3.5}}}',
                '{"node_name": "n", "state": []}', '[1]', '{"node_name": "n", "state": {"flight": {"armed": "yes"}}}'):
        try:
            s.report_decode(bad)
        except RuntimeError as e:
            assert "JSON" in str(e)
        else:
            raise AssertionError(f"accepted {bad!r}")
    s.close()


def test_to_lower_matches_go_for_common_scripts(pkg):
    """strings.ToLower (controller.go:177, :188, :206) is Unicode-aware: a preferred node written
    'ÉDGE-1' or 'УЗЕЛ-2' must still match. Python's str.lower agrees with Go's unicode.ToLower on
    these ranges except U+0130, which Go maps to plain 'i'."""
    tl = pkg.hostbinding.to_lower
    for s in ("Node-1", "ÉDGE-Ünïté", "ÀÁÂÃÄÅÆÇÈÉÊËÌÍÎÏÐÑÒÓÔÕÖ×ØÙÚÛÜÝÞß", "ĀĂĄĆĈĊČĎĐĒĔĖĘĚĜĞĠĢĤĦĨĪĬĮĲĴĶĹĻĽĿŁŃŅŇŊŌŎŐŒŔŖŘŚŜŞŠŢŤŦŨŪŬŮŰŲŴŶŸŹŻŽ",
              "ΑΒΓΔΕΖΗΘΙΚΛΜΝΞΟΠΡΣΤΥΦΧΨΩΆΈΉΊΌΎΏ", "УЗЕЛ-2 ЁЂЃЄЅІЇЈЉЊЋЌЍЎЏ АБВГДЕЖЗИЙКЛМНОПРСТУФХЦЧШЩЪЫЬЭЮЯ", "日本語-Node", "mixed-ß-ſ"):
        assert tl(s) == s.lower(), s
    assert tl("İstanbul") == "istanbul"                       # Go: U+0130 -> 'i'
    assert tl(b"N\xffode-\xc9") == "n�ode-�"         # invalid bytes in a non-ASCII string: U+FFFD, like strings.Map
    assert tl(b"ASCII\x7f") == "ascii\x7f"


def test_status_write_back_is_concurrent_and_complete(pkg):
    """WriteStatuses (controller.go:223-250 batched): every updated status is encoded and handed to
    the injected UpdateStatus exactly once, from several threads; skipped requests are not written;
    failures are reported per request."""
    import threading
    s = pkg.hostbinding.HostSession(None)
    n = 200
    for i in range(n):
        s.add_status(i % 7 != 0, "Assigned" if i % 3 else "", f"node-{i}", f"uav-{i}", 90.0 + (i % 5) * 2.5,
                     "选中节点 node-%d (电量 %.1f%%)" % (i, 80.0))
    seen, threads, lock = {}, set(), threading.Lock()

    def update(ns, name, js):
        with lock:
            assert name not in seen
            seen[name] = json.loads(js)
            threads.add(threading.get_ident())
        return int(name.split("-")[1]) % 50 != 49            # a few refused updates
    failed = s.write_statuses("2026-01-01T00:00:00Z", 8, update)
    want = [i for i in range(n) if i % 7 != 0]
    assert sorted(int(k.split("-")[1]) for k in seen) == want
    assert failed == sum(1 for i in want if i % 50 == 49)
    for i in want:
        st = seen[f"req-{i}"]
        assert st["phase"] == ("Assigned" if i % 3 else "Pending")          # controller.go:238-240
        assert st["assignedNode"] == f"node-{i}" and st["score"] == 90.0 + (i % 5) * 2.5
        assert st["lastUpdated"] == "2026-01-01T00:00:00Z" and st["message"].startswith("选中节点")
    s.close()
