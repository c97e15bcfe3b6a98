"""Mechanical check of the cgo sources against include/nosplan.h (they cannot be compiled here: no Go toolchain).
  * every C.nosplan_<function>(...) call must name a function declared in the header, with the same number of arguments;
  * every field named in a C.nosplan_<struct>{ field: ... } composite literal must be a field of that struct;
  * every C.NOSPLAN_<CONSTANT> must be #defined in the header.
Exit code 1 and one line per problem otherwise.  Run by tests/test_go_binding.py."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def strip_comments(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return re.sub(r"//[^\n]*", " ", text)


def split_args(s):
    depth, cur, out = 0, "", []
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur); cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return out


def balanced(text, start, open_ch, close_ch):
    depth, i = 0, start
    while i < len(text):
        if text[i] == open_ch:
            depth += 1
        elif text[i] == close_ch:
            depth -= 1
            if depth == 0:
                return text[start + 1:i]
        i += 1
    raise ValueError("unbalanced")


def header_model(path):
    h = strip_comments(open(path).read())
    funcs = {}
    for m in re.finditer(r"\b(?:int|void|int32_t|int64_t|const char\s*\*|void\s*\*)\s*(nosplan_\w+)\s*\(", h):
        args = balanced(h, m.end() - 1, "(", ")")
        n = 0 if args.strip() in ("", "void") else len(split_args(args))
        funcs[m.group(1)] = n
    structs = {}
    for m in re.finditer(r"typedef\s+struct\s+(nosplan_\w+)\s*\{", h):
        body = balanced(h, m.end() - 1, "{", "}")
        fields = set()
        for decl in body.split(";"):
            for name in re.findall(r"(\w+)\s*(?:,|$|\[)", decl.strip().replace("*", " ")):
                fields.add(name)
            mm = re.search(r"(\w+)\s*$", decl.strip())
            if mm:
                fields.add(mm.group(1))
            for part in decl.split(","):
                mm = re.search(r"(\w+)\s*$", part.strip().replace("*", " "))
                if mm:
                    fields.add(mm.group(1))
        structs[m.group(1)] = fields
    consts = set(re.findall(r"#define\s+(NOSPLAN_\w+)", h))
    return funcs, structs, consts


def check(go_path, funcs, structs, consts):
    src = strip_comments(open(go_path).read())
    problems, n_calls = [], 0
    for m in re.finditer(r"\bC\.(nosplan_\w+)\s*\(", src):
        name = m.group(1)
        if name in structs:          # a conversion, not a call
            continue
        n_calls += 1
        args = split_args(balanced(src, m.end() - 1, "(", ")"))
        if name not in funcs:
            problems.append(f"{go_path}: C.{name} is not declared in nosplan.h")
        elif len(args) != funcs[name]:
            problems.append(f"{go_path}: C.{name} called with {len(args)} arguments, nosplan.h declares {funcs[name]}")
    for m in re.finditer(r"\bC\.(nosplan_\w+)\s*\{", src):
        name = m.group(1)
        if name not in structs:
            problems.append(f"{go_path}: C.{name} is not a struct of nosplan.h")
            continue
        for part in split_args(balanced(src, m.end() - 1, "{", "}")):
            mm = re.match(r"\s*(\w+)\s*:", part)
            if mm and mm.group(1) not in structs[name]:
                problems.append(f"{go_path}: {name} has no field {mm.group(1)}")
    for c in set(re.findall(r"\bC\.(NOSPLAN_\w+)", src)):
        if c not in consts:
            problems.append(f"{go_path}: C.{c} is not defined in nosplan.h")
    return problems, n_calls


def main():
    funcs, structs, consts = header_model(os.path.join(ROOT, "include", "nosplan.h"))
    problems, total = [], 0
    for f in sorted(os.listdir(os.path.dirname(os.path.abspath(__file__)))):
        if f.endswith(".go"):
            p, n = check(os.path.join(os.path.dirname(os.path.abspath(__file__)), f), funcs, structs, consts)
            problems += p
            total += n
    for p in problems:
        print(p)
    print(f"{total} C calls checked against {len(funcs)} declarations, {len(structs)} structs, {len(consts)} constants: "
          f"{'OK' if not problems else str(len(problems)) + ' problem(s)'}")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main())
