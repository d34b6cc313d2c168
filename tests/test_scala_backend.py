"""The Scala back-end overlay (Compiler/) against the C ABI and against the reference tree, as far as text goes: no JVM
exists here, so the overlay cannot be compiled; these checks keep it honest about names.

1. ExaMG_Symbols (every examg_* name the back-end can print) is a subset of include/examg.h, and covers every hot-path
   entry point of the header.
2. Every examg_* string literal anywhere in the overlay is declared in the header.
3. Every reference class the overlay matches on / constructs exists in the reference with at least the number of
   constructor parameters used (against /root/reference when present; the extracted table is committed under
   tests/golden/ so the check also runs on machines without the reference).
"""
import json
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OVERLAY = os.path.join(ROOT, "Compiler", "src", "exastencils")
REF = "/root/reference/Compiler/src/exastencils"
TABLE = os.path.join(ROOT, "tests", "golden", "scala_reference_classes.json")


def _overlay_sources():
    out = {}
    for dirpath, _, files in os.walk(OVERLAY):
        for fn in files:
            if fn.endswith(".scala"):
                out[os.path.join(dirpath, fn)] = open(os.path.join(dirpath, fn)).read()
    return out


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "examg.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(examg_[a-z0-9_]+)\s*\(", text))


def test_symbol_table_matches_header():
    src = open(os.path.join(OVERLAY, "parallelization", "api", "examg", "ExaMG_Call.scala")).read()
    block = src[src.index("object ExaMG_Symbols"):src.index("/// ExaMG_Context")]
    table = set(re.findall(r'val \w+ = "(examg_[a-z0-9_]+)"', block))
    header = _header_symbols()
    assert len(table) >= 55
    assert not (table - header), "printed by the back-end but not in examg.h: %s" % sorted(table - header)
    # every loop-shaped / per-statement entry point of the header has to be reachable from the generator
    hot = {s for s in header if re.match(
        r"examg_(apply|residual|residual_restrict|smooth_\w+|restrict\w*|prolong_add\w*|dot|max_abs_diff|set|copy|axpy|xpay|"
        r"exchange|apply_bc|cg_coarse|bicgstab_coarse|cr_coarse|minres_coarse|smoother_coarse|cycle_desc_from_options|system2_\w+|field_(create\w*|set_bc\w*|eval|"
        r"advance|upload|download|print|write|read|fill)|stencil_\w+|init|finalize|comm_(export|import|blob_size)|timer_(start|stop|print_all)|benchmark_(start|stop))$", s)}
    assert not (hot - table), "hot-path entry points the back-end never prints: %s" % sorted(hot - table)


def test_every_literal_is_declared():
    header = _header_symbols()
    for path, text in _overlay_sources().items():
        code = re.sub(r"//.*", "", text)
        types = {"examg_ctx", "examg_config", "examg_bc_op"}
        for name in set(re.findall(r'"(examg_[a-z0-9_]+)"', code)) - types:
            assert name in header, "%s prints %s, which examg.h does not declare" % (os.path.relpath(path, ROOT), name)


# reference classes used by the overlay: name -> minimal number of constructor parameters relied on
USED = {
    "IR_LoopOverPoints": 10, "IR_Communicate": 6, "IR_ApplyBC": 2, "IR_AdvanceSlot": 2, "IR_SlotAccess": 2,
    "IR_IV_ActiveSlot": 2, "IR_Assignment": 3, "IR_Multiplication": 1, "IR_Addition": 1, "IR_Subtraction": 2,
    "IR_Division": 2, "IR_Modulo": 2, "IR_EqEq": 2, "IR_AndAnd": 2, "IR_Negative": 1, "IR_Power": 2, "IR_FunctionCall": 2,
    "IR_StencilAccess": 2, "IR_StencilFieldAccess": 5, "IR_Stencil": 5, "IR_StencilField": 4,     "IR_Field": 9, "IR_LocalSolve": 4, "IR_DirichletBC": 2, "IR_NeumannBC": 1, "IR_PrintField": 8, "IR_WriteField": 8, "IR_ReadField": 8,
    "IR_CopyFromExternalField": 2, "IR_CopyToExternalField": 2, "IR_VariableDeclaration": 3, "IR_IfCondition": 3,
    "IR_WhileLoop": 2, "IR_Scope": 1, "IR_AllocateDataFunction": 2,
    "IR_GlobalCollection": 1, "IR_RealConstant": 1, "IR_IntegerConstant": 1, "IR_StringConstant": 1,
    "IR_PointerDatatype": 1, "IR_SpecialDatatype": 1, "IR_DomainFromAABB": 2,
    "ConditionedSingleStrategyWrapper": 2, "IR_AtFaceCenter": 1,
}
OBJECTS = ["IR_ResolveLoopOverPoints", "IR_SetupAllocateDataFunctionWrapper", "IR_ResolveBenchmarkFunctions",
           "IR_DefaultLayerHandler", "IR_FieldCollection", "IR_StencilCollection", "IR_StencilFieldCollection",
           "IR_DomainCollection", "IR_NoBC", "IR_NullStatement", "IR_AtCellCenter", "MPI_IV_MpiRank", "IR_FieldIteratorAccess"]


def _count_params(sig):
    depth, n, cur = 0, 0, ""
    for ch in sig:
        if ch in "([":
            depth += 1
        elif ch in ")]":
            depth -= 1
        if ch == "," and depth == 0:
            n += 1
            cur = ""
        else:
            cur += ch
    return n + (1 if cur.strip() else 0)


def _extract_reference_table():
    classes, objects = {}, set()
    for dirpath, _, files in os.walk(REF):
        for fn in files:
            if not fn.endswith(".scala"):
                continue
            text = open(os.path.join(dirpath, fn), errors="replace").read()
            for m in re.finditer(r"^case class (\w+)\s*\(", text, flags=re.M):
                name = m.group(1)
                if name not in USED:
                    continue
                i, depth = m.end(), 1
                while depth and i < len(text):
                    depth += text[i] in "([" 
                    depth -= text[i] in ")]"
                    i += 1
                classes[name] = _count_params(text[m.end():i - 1])
            for m in re.finditer(r"^(?:case )?object (\w+)", text, flags=re.M):
                if m.group(1) in OBJECTS:
                    objects.add(m.group(1))
    return {"classes": classes, "objects": sorted(objects)}


def test_reference_classes_exist():
    if os.path.isdir(REF):
        table = _extract_reference_table()
        if not os.path.exists(TABLE) or json.load(open(TABLE)) != table:
            with open(TABLE, "w") as fh:
                json.dump(table, fh, indent=1, sort_keys=True)
    table = json.load(open(TABLE))
    missing = [c for c in USED if c not in table["classes"]]
    assert not missing, "not in the reference: %s" % missing
    short = {c: (table["classes"][c], n) for c, n in USED.items() if table["classes"][c] < n}
    assert not short, "constructor has fewer parameters than the overlay uses: %s" % short
    assert not (set(OBJECTS) - set(table["objects"])), sorted(set(OBJECTS) - set(table["objects"]))
    # and the overlay really refers to them
    text = "\n".join(_overlay_sources().values())
    unused = [c for c in list(USED) + OBJECTS if c not in text]
    assert not unused, "listed but not used by the overlay: %s" % unused


def test_patch_touches_only_known_lines():
    patch = open(os.path.join(ROOT, "Compiler", "patches", "0001-knowledge-flag-and-handler-hook.patch")).read()
    assert "experimental_examg" in patch and "IR_ExaMGLayerHandler" in patch
    if os.path.isdir(REF):
        ctx = {"config/Knowledge.scala": "var experimental_splitLoopsForAsyncComm : Boolean = false",
               "app/ExaLayerHandler.scala": "ir_handler = IR_WaLBerlaLayerHandler",
               "prettyprinting/MakefileGenerator.scala": "Platform.resolveLdFlags"}
        for rel, line in ctx.items():
            assert line in open(os.path.join(REF, rel)).read(), rel
