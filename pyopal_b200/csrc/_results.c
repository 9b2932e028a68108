/* pyopal_b200/csrc/_results.c -- C storage of the search results and the bulk
 * constructor of the result list.
 *
 * The reference embeds an OpalSearchResult in every result object
 * (/root/reference/src/pyopal/lib.pxd:113-126) and its plugin creates the N
 * objects of a search in a tight Cython loop (platform/pyx.in:65-72).  With the
 * DP on the GPU that loop is what a caller of Aligner.align waits for (0.5 s
 * for 570k targets in pure Python against a 13 ms search), so the storage and
 * the loop are native here too: `ResultBase` holds the fields, the Python
 * classes ScoreResult / EndResult / FullResult (lib.py) derive from it, and
 * `make_results` fills a list straight from the arrays the C-ABI returned.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stddef.h>
#include <stdint.h>
#include <structmember.h>

typedef struct {
    PyObject_HEAD
    int target_index; /* defaults of __cinit__, lib.pyx:787-795 */
    int score;
    int score_set;
    int query_end, target_end;
    int query_start, target_start;
    int query_length, target_length;
    PyObject* alignment; /* bytes of operations (M=0 D=1 I=2 X=3) or NULL */
} ResultBase;

static void result_init_fields(ResultBase* r) {
    r->target_index = -1;
    r->score = 0;
    r->score_set = 0;
    r->query_end = r->target_end = -1;
    r->query_start = r->target_start = -1;
    r->query_length = r->target_length = -1;
    r->alignment = NULL;
}

static PyObject* result_new(PyTypeObject* type, PyObject* args, PyObject* kwds) {
    (void)args; (void)kwds;
    ResultBase* r = (ResultBase*)type->tp_alloc(type, 0);
    if (r) result_init_fields(r);
    return (PyObject*)r;
}

static void result_dealloc(ResultBase* r) {
    Py_XDECREF(r->alignment);
    Py_TYPE(r)->tp_free((PyObject*)r);
}

static PyObject* result_get_score(ResultBase* r, void* closure) {
    (void)closure;
    if (!r->score_set) Py_RETURN_NONE;
    return PyLong_FromLong(r->score);
}

static int result_set_score(ResultBase* r, PyObject* value, void* closure) {
    (void)closure;
    if (value == NULL || value == Py_None) { r->score_set = 0; r->score = 0; return 0; }
    long v = PyLong_AsLong(value);
    if (v == -1 && PyErr_Occurred()) return -1;
    if (v < INT32_MIN || v > INT32_MAX) { PyErr_SetString(PyExc_OverflowError, "score out of range"); return -1; }
    r->score = (int)v;
    r->score_set = 1;
    return 0;
}

static PyMemberDef result_members[] = {
    {"_target_index", T_INT, offsetof(ResultBase, target_index), 0, NULL},
    {"_query_end", T_INT, offsetof(ResultBase, query_end), 0, NULL},
    {"_target_end", T_INT, offsetof(ResultBase, target_end), 0, NULL},
    {"_query_start", T_INT, offsetof(ResultBase, query_start), 0, NULL},
    {"_target_start", T_INT, offsetof(ResultBase, target_start), 0, NULL},
    {"_query_length", T_INT, offsetof(ResultBase, query_length), 0, NULL},
    {"_target_length", T_INT, offsetof(ResultBase, target_length), 0, NULL},
    {"_alignment", T_OBJECT, offsetof(ResultBase, alignment), 0, NULL},
    {NULL, 0, 0, 0, NULL}};

static PyGetSetDef result_getset[] = {
    {"_score", (getter)result_get_score, (setter)result_set_score, NULL, NULL},
    {NULL, NULL, NULL, NULL, NULL}};

static PyTypeObject ResultBaseType = {
    PyVarObject_HEAD_INIT(NULL, 0).tp_name = "pyopal_b200._results.ResultBase",
    .tp_basicsize = sizeof(ResultBase),
    .tp_dealloc = (destructor)result_dealloc,
    .tp_flags = Py_TPFLAGS_DEFAULT | Py_TPFLAGS_BASETYPE,
    .tp_doc = "C storage of one search result (the OpalSearchResult of opal.pxd:26-34).",
    .tp_members = result_members,
    .tp_getset = result_getset,
    .tp_new = result_new,
};

/* ---- make_results ---------------------------------------------------------
 * make_results(cls, mode, start, n, scores, end_q, end_t, start_q, start_t,
 *              query_length, lengths, ops, ops_offsets, ops_len) -> list
 * mode 0 = score, 1 = end, 2 = full (OPAL_SEARCH_*).  Arrays are buffers of
 * int32 (`lengths` holds the whole database, indexed start + i), `ops` a byte
 * buffer, `ops_offsets` int64; unused ones are None. */
typedef struct {
    Py_buffer view;
    int have;
} Buf;

static int buf_get(PyObject* obj, Buf* b, Py_ssize_t itemsize, Py_ssize_t need, const char* name) {
    b->have = 0;
    if (obj == Py_None) return 0;
    if (PyObject_GetBuffer(obj, &b->view, PyBUF_SIMPLE) < 0) return -1;
    b->have = 1;
    if (b->view.len < need * itemsize) {
        PyErr_Format(PyExc_ValueError, "buffer `%s` is too small", name);
        return -1;
    }
    return 0;
}

static PyObject* make_results(PyObject* self, PyObject* args) {
    (void)self;
    PyObject *cls, *o_sc, *o_eq, *o_et, *o_sq, *o_st, *o_len, *o_ops, *o_oo, *o_ol;
    int mode, qlen;
    Py_ssize_t start, n;
    if (!PyArg_ParseTuple(args, "OinnOOOOOiOOOO", &cls, &mode, &start, &n, &o_sc, &o_eq, &o_et, &o_sq, &o_st, &qlen,
                          &o_len, &o_ops, &o_oo, &o_ol))
        return NULL;
    if (!PyType_Check(cls) || !PyType_IsSubtype((PyTypeObject*)cls, &ResultBaseType)) {
        PyErr_SetString(PyExc_TypeError, "cls must derive from ResultBase");
        return NULL;
    }
    if (n < 0 || start < 0 || mode < 0 || mode > 2) {
        PyErr_SetString(PyExc_ValueError, "invalid arguments");
        return NULL;
    }
    PyTypeObject* type = (PyTypeObject*)cls;
    Buf sc, eq, et, sq, st, ln, ops, oo, ol;
    Buf* all[9] = {&sc, &eq, &et, &sq, &st, &ln, &ops, &oo, &ol};
    for (int i = 0; i < 9; i++) all[i]->have = 0;
    PyObject* list = NULL;
    int bad = 0;
    bad |= buf_get(o_sc, &sc, 4, n, "scores") < 0;
    if (!bad && mode >= 1) bad |= buf_get(o_eq, &eq, 4, n, "end_q") < 0 || buf_get(o_et, &et, 4, n, "end_t") < 0;
    if (!bad && mode == 2)
        bad |= buf_get(o_sq, &sq, 4, n, "start_q") < 0 || buf_get(o_st, &st, 4, n, "start_t") < 0 ||
               buf_get(o_len, &ln, 4, start + n, "lengths") < 0 || buf_get(o_ops, &ops, 1, 0, "ops") < 0 ||
               buf_get(o_oo, &oo, 8, n, "ops_offsets") < 0 || buf_get(o_ol, &ol, 4, n, "ops_len") < 0;
    if (!bad && (!sc.have || (mode >= 1 && (!eq.have || !et.have)) ||
                 (mode == 2 && (!sq.have || !st.have || !ln.have || !ops.have || !oo.have || !ol.have)))) {
        PyErr_SetString(PyExc_ValueError, "missing result array");
        bad = 1;
    }
    if (!bad) list = PyList_New(n);
    if (list) {
        const int32_t* p_sc = (const int32_t*)sc.view.buf;
        const int32_t* p_eq = eq.have ? (const int32_t*)eq.view.buf : NULL;
        const int32_t* p_et = et.have ? (const int32_t*)et.view.buf : NULL;
        const int32_t* p_sq = sq.have ? (const int32_t*)sq.view.buf : NULL;
        const int32_t* p_st = st.have ? (const int32_t*)st.view.buf : NULL;
        const int32_t* p_ln = ln.have ? (const int32_t*)ln.view.buf : NULL;
        const int64_t* p_oo = oo.have ? (const int64_t*)oo.view.buf : NULL;
        const int32_t* p_ol = ol.have ? (const int32_t*)ol.view.buf : NULL;
        const char* p_ops = ops.have ? (const char*)ops.view.buf : NULL;
        for (Py_ssize_t i = 0; i < n; i++) {
            ResultBase* r = (ResultBase*)type->tp_alloc(type, 0);
            if (!r) { Py_CLEAR(list); break; }
            result_init_fields(r);
            PyList_SET_ITEM(list, i, (PyObject*)r);
            r->target_index = (int)(start + i); /* absolute index, pyx.in:69 */
            r->score = p_sc[i];
            r->score_set = 1;
            if (mode >= 1) { r->query_end = p_eq[i]; r->target_end = p_et[i]; }
            if (mode == 2) {
                r->query_start = p_sq[i];
                r->target_start = p_st[i];
                r->query_length = qlen; /* pyx.in:95-99 */
                r->target_length = p_ln[start + i];
                const int64_t off = p_oo[i];
                const int32_t len = p_ol[i];
                if (len < 0 || off < 0 || off + len > ops.view.len) {
                    PyErr_SetString(PyExc_ValueError, "alignment slice out of range");
                    Py_CLEAR(list);
                    break;
                }
                r->alignment = PyBytes_FromStringAndSize(p_ops + off, len);
                if (!r->alignment) { Py_CLEAR(list); break; }
            }
        }
    }
    for (int i = 0; i < 9; i++)
        if (all[i]->have) PyBuffer_Release(&all[i]->view);
    return list;
}

static PyMethodDef module_methods[] = {
    {"make_results", make_results, METH_VARARGS, "Build the list of result objects of one search from the C-ABI's arrays."},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moduledef = {PyModuleDef_HEAD_INIT, "_results", "Native result storage of pyopal_b200.", -1,
                                       module_methods, NULL, NULL, NULL, NULL};

PyMODINIT_FUNC PyInit__results(void) {
    if (PyType_Ready(&ResultBaseType) < 0) return NULL;
    PyObject* m = PyModule_Create(&moduledef);
    if (!m) return NULL;
    Py_INCREF(&ResultBaseType);
    if (PyModule_AddObject(m, "ResultBase", (PyObject*)&ResultBaseType) < 0) {
        Py_DECREF(&ResultBaseType);
        Py_DECREF(m);
        return NULL;
    }
    return m;
}
