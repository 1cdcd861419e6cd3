"""CPU fuzz of the wire decoder (host/klm_wire.cpp, a flat-array JSON document since round 2)
against Python's json + the string-level oracle's coercions: random documents with missing and
mistyped fields, nested junk, duplicate keys (the last one wins, as in a Go map), escapes and
non-ASCII text, numbers in every JSON spelling, random whitespace."""
import json
import random

import controller_oracle as co


def _num_text(rng, v):
    """One of several JSON spellings of the number v (all parse to the same float64)."""
    if isinstance(v, int):
        return rng.choice([str(v), f"{v}.0", f"{v}e0", f"{v}E+0", f"{v * 10}e-1"]) if abs(v) < 10**6 else str(v)
    return rng.choice([repr(v), f"{v:.17g}", f"{v:.17e}".replace("e+0", "E+").replace("e-0", "e-") if v else "0.0"])


def _junk(rng, depth=0):
    k = rng.randrange(8 if depth < 3 else 5)
    if k == 0: return None
    if k == 1: return rng.random() < 0.5
    if k == 2: return rng.randrange(-10**9, 10**9)
    if k == 3: return rng.random() * 10 ** rng.randrange(-5, 6)
    if k == 4: return "".join(rng.choice('ab"\\/\n\téЖ中\U0001f600 {}[],:') for _ in range(rng.randrange(0, 9)))
    if k == 5: return [_junk(rng, depth + 1) for _ in range(rng.randrange(0, 4))]
    return {f"k{rng.randrange(5)}": _junk(rng, depth + 1) for _ in range(rng.randrange(0, 4))}


def _dump(rng, obj):
    """json.dumps with random escaping / spacing, then random extra whitespace between tokens."""
    text = json.dumps(obj, ensure_ascii=rng.random() < 0.5,
                      separators=rng.choice([(",", ":"), (", ", ": "), (" ,\n", " :\t")]))
    return rng.choice(["", " ", "\n\t"]) + text + rng.choice(["", " ", "\r\n"])


def test_uav_list_decode_fuzz(pkg):
    rng = random.Random(1234)
    s = pkg.hostbinding.HostSession(None)
    for _ in range(300):
        items = []
        for _ in range(rng.randrange(0, 12)):
            it = {}
            if rng.random() < 0.9:
                spec = {}
                if rng.random() < 0.85:
                    spec["node_name"] = rng.choice(["Node-1", "", "nÉЖ", 'q"uo\\te', 7, None, ["x"]])
                if rng.random() < 0.7:
                    spec["uav_id"] = rng.choice(["uav-1", 3, ""])
                if rng.random() < 0.85:
                    spec["battery"] = rng.choice([{"remaining_percent": rng.choice([90, 85.5, 0, -1, 1e2, "90", None, True])},
                                                  {}, 5, None])
                for _ in range(rng.randrange(0, 3)):
                    spec[f"x{rng.randrange(9)}"] = _junk(rng)
                it["spec"] = rng.choice([spec, spec, spec, 5, None])
            if rng.random() < 0.8:
                it["status"] = rng.choice([{"collection_status": rng.choice(["active", "ACTIVE", "stale", "", 3, None])},
                                           {}, "x"])
            for _ in range(rng.randrange(0, 2)):
                it[f"y{rng.randrange(9)}"] = _junk(rng)
            items.append(it)
        doc = {"items": items} if rng.random() < 0.7 else items
        if isinstance(doc, dict) and rng.random() < 0.3:
            doc = {"kind": "UAVMetricList", "metadata": _junk(rng), **doc}
        text = _dump(rng, doc)
        assert s.load_uavs_json(text) == len(items), text
        for i, it in enumerate(items):
            u = s.uav_info(i)
            spec = it.get("spec") if isinstance(it.get("spec"), dict) else {}
            status = it.get("status") if isinstance(it.get("status"), dict) else {}
            nn = spec.get("node_name")
            assert u["has_name"] == isinstance(nn, str) and u["name"] == (nn if isinstance(nn, str) else ""), (text, i)
            assert u["battery"] == co.read_float(spec, "battery", "remaining_percent"), (text, i)
            cs = status.get("collection_status")
            assert u["has_status"] == isinstance(cs, str) and u["status"] == (cs if isinstance(cs, str) else ""), (text, i)
    s.close()


def test_request_list_decode_fuzz(pkg):
    rng = random.Random(99)
    s = pkg.hostbinding.HostSession(None)
    for _ in range(300):
        specs, items = [], []
        for _ in range(rng.randrange(0, 10)):
            sp = {}
            if rng.random() < 0.9:
                sp["workload"] = rng.choice([{"name": "w", "namespace": "d"}, {"name": "jüb", "namespace": "n", "type": "t"},
                                             {"name": 5}, {}, "w", None])
            if rng.random() < 0.8:
                sp["minBatteryPercent"] = rng.choice([80, 80.0, 8e1, 0, 50.5, -3, "80", None, True, 10**15])
            if rng.random() < 0.8:
                sp["preferredNodes"] = rng.choice([["a", "B"], [], ["a", 5, None, "c", ["d"]], "a", None,
                                                   ["nÉ", 'q"\\']])
            for _ in range(rng.randrange(0, 2)):
                sp[f"z{rng.randrange(9)}"] = _junk(rng)
            it = {"spec": sp}
            if rng.random() < 0.6:
                it["status"] = rng.choice([{"phase": rng.choice(["Pending", "Assigned", "", 4])}, {}, None])
            specs.append(sp)
            items.append(it)
        text = _dump(rng, {"items": items})
        for quirk in (True, False):
            assert s.load_requests_json(text, apimachinery_ints=quirk) == len(items), text
            for i, (sp, it) in enumerate(zip(specs, items)):
                want = co.decode_spec(sp, apimachinery_ints=quirk)
                got = s.request_info(i)
                assert got["min_battery"] == want.min_battery_percent, (text, i, quirk)
                assert got["n_pref"] == len(want.preferred_nodes), (text, i)
                assert got["workload"] == want.workload.name, (text, i)
    s.close()


def test_number_spellings_and_duplicate_keys(pkg):
    rng = random.Random(5)
    s = pkg.hostbinding.HostSession(None)
    for _ in range(400):
        v = rng.choice([rng.randrange(0, 101), round(rng.random() * 100, rng.randrange(0, 8)), 1e-7, 123456789012, 0.1 + 0.2])
        first = rng.choice(["1", '"x"', "null", "{}"])
        text = ('{"items":[{"spec":{"battery":{"remaining_percent":%s,"remaining_percent":%s},"node_name":"a","node_name":"b"}}]}'
                % (first, _num_text(rng, v)))
        assert s.load_uavs_json(text) == 1
        u = s.uav_info(0)
        assert u["battery"] == float(v), (text, u)
        assert u["name"] == "b"
    s.close()
