"""WebUI page: the script is structurally well formed, every REST path it calls is a route the master registers, and the
master serves it at /ui (reference: webui/react — Dashboard, ExperimentList, ExperimentDetails, TrialDetails, TaskList,
Cluster, MasterLogs pages)."""
import pathlib
import re

ROOT = pathlib.Path(__file__).resolve().parent.parent
PAGE = ROOT / "determined_b200" / "master" / "webui" / "index.html"


def _script() -> str:
    html = PAGE.read_text()
    return html[html.index("<script>") + 8:html.rindex("</script>")]


def _check_balanced(src: str) -> None:
    """A small JS scanner: strings, template literals (with nested ${...}), regex literals and comments are skipped;
    (), [], {} must nest properly."""
    stack = []
    i, n = 0, len(src)
    prev_sig = ""          # last significant character, to tell division from a regex literal
    close = {")": "(", "]": "[", "}": "{"}
    while i < n:
        c = src[i]
        if c in " \t\r\n":
            i += 1
            continue
        if src.startswith("//", i):
            i = src.index("\n", i) if "\n" in src[i:] else n
            continue
        if src.startswith("/*", i):
            i = src.index("*/", i) + 2
            continue
        if c in "'\"":
            j = i + 1
            while src[j] != c:
                j += 2 if src[j] == "\\" else 1
                assert src[j - 1] != "\n" or src[j - 2] == "\\", f"unterminated string at offset {i}: {src[i:i + 60]!r}"
            i = j + 1
            prev_sig = "x"
            continue
        if c == "`":
            stack.append("`")
            i += 1
            # scan template text until ${ or closing backtick
            while True:
                if src[i] == "\\":
                    i += 2
                elif src[i] == "`":
                    assert stack.pop() == "`"
                    i += 1
                    break
                elif src.startswith("${", i):
                    stack.append("${")
                    i += 2
                    break
                else:
                    i += 1
            prev_sig = "x"
            continue
        if c == "/" and prev_sig in "(,=:[!&|?{};\n" + "":
            j = i + 1
            in_class = False
            while src[j] != "/" or in_class:
                if src[j] == "\\":
                    j += 1
                elif src[j] == "[":
                    in_class = True
                elif src[j] == "]":
                    in_class = False
                j += 1
            i = j + 1
            while i < n and src[i].isalpha():
                i += 1
            prev_sig = "x"
            continue
        if c in "([{":
            stack.append(c)
        elif c in ")]}":
            assert stack, f"unbalanced {c!r} at offset {i}: {src[max(0, i - 60):i + 20]!r}"
            top = stack.pop()
            if c == "}" and top == "${":
                # back inside the template literal
                while True:
                    if src[i + 1] == "\\":
                        i += 2
                    elif src[i + 1] == "`":
                        assert stack.pop() == "`"
                        i += 1
                        break
                    elif src.startswith("${", i + 1):
                        stack.append("${")
                        i += 2
                        break
                    else:
                        i += 1
            else:
                assert top == close[c], f"mismatched {top!r} ... {c!r} at offset {i}: {src[max(0, i - 80):i + 20]!r}"
        prev_sig = c
        i += 1
    assert not stack, f"unclosed {stack[-5:]}"


def test_script_is_well_formed():
    _check_balanced(_script())


def test_every_api_path_is_a_master_route():
    api_src = (ROOT / "csrc" / "core" / "master_api.cpp").read_text()
    routes = set(re.findall(r'"(/api/v1/[^"]*)"', api_src))
    routes |= {"/api/v1/" + k for k in ("commands", "notebooks", "shells", "tensorboards")}          # registered in a loop over kinds
    routes |= {"/api/v1/" + k + "/:id/kill" for k in ("commands", "notebooks", "shells", "tensorboards")}
    routes |= {"/api/v1/experiments/:id/" + v for v in ("activate", "pause", "cancel", "kill", "archive", "unarchive")}
    patterns = [re.compile("^" + re.sub(r":\w+", "[^/]+", r) + "$") for r in routes]
    called = set()
    for m in re.finditer(r'[`"](/api/v1/[^`"?]*)', _script()):
        path = re.sub(r"\$\{[^}]*\}", "X", m.group(1))
        called.add(path)
    assert len(called) >= 15
    # templated over the four command kinds / over the experiment and trial verbs (all registered, see `routes` above)
    called -= {"/api/v1/X/X/kill", "/api/v1/", "/api/v1/experiments/X/X", "/api/v1/trials/X/X"}
    for path in sorted(called):
        assert any(p.match(path) for p in patterns), f"the page calls {path}, which the master does not serve"


def test_pages_present():
    s = _script()
    for page in ("dashboard", "experiments", "experiment", "trial", "tasks", "cluster", "logs", "login"):
        assert re.search(r"\b%s\(" % page, s), page


def test_rest_api_doc_is_current():
    import subprocess
    import sys

    assert subprocess.run([sys.executable, str(ROOT / "scripts" / "gen_rest_api_doc.py"), "--check"]).returncode == 0, \
        "docs/rest-api.md is stale: run python scripts/gen_rest_api_doc.py"
