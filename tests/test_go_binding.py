"""integration/go/*.go cannot be compiled here (no Go toolchain): every C.nosplan_* call, struct literal and constant in
them is checked mechanically against include/nosplan.h, and the checker itself is checked on a broken copy."""
import importlib.util
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("check_arity", os.path.join(ROOT, "integration", "go", "check_arity.py"))
ca = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ca)


def test_go_sources_match_the_header():
    funcs, structs, consts = ca.header_model(os.path.join(ROOT, "include", "nosplan.h"))
    assert funcs["nosplan_plan"] == 7 and funcs["nosplan_create"] == 2 and "gpu_used_mask" in structs["nosplan_snapshot"]
    total = 0
    for f in ("planner_cgo.go", "capacityscheduling_cgo.go"):
        problems, n = ca.check(os.path.join(ROOT, "integration", "go", f), funcs, structs, consts)
        assert not problems, problems
        total += n
    assert total >= 10


def test_the_checker_sees_a_wrong_call(tmp_path):
    funcs, structs, consts = ca.header_model(os.path.join(ROOT, "include", "nosplan.h"))
    src = open(os.path.join(ROOT, "integration", "go", "planner_cgo.go")).read()
    bad = src.replace("C.nosplan_plan(p.h, &snap, &cp, pu8(pre),", "C.nosplan_plan(p.h, &snap, &cp,").replace("gpu_free_mask:", "gpu_freee_mask:") \
             .replace("C.NOSPLAN_FILTER_RESOURCES_FIT", "C.NOSPLAN_FILTER_RESOURCE_FIT")
    p = tmp_path / "bad.go"
    p.write_text(bad)
    problems, _ = ca.check(str(p), funcs, structs, consts)
    text = "\n".join(problems)
    assert "called with 6 arguments" in text and "gpu_freee_mask" in text and "NOSPLAN_FILTER_RESOURCE_FIT" in text
