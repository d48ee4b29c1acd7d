/* _evi_records — CPython helper that turns the [B, max_k] host columns of one batch's selected triples into the
 * reference's per-question `triplets_by_k` dictionaries (src/callbacks/retriever_topk_edge_writer.py:321-349):
 *
 *     {k: [{"head_entity_id", "relation_id", "tail_entity_id", "head_text", "relation_text", "tail_text", "score",
 *           "label", "rank"[, "logit_fwd"][, "logit_bwd"]}, ... min(k, count) entries] for k in topk_values}
 *
 * The reference builds ~100 dicts per question in Python; at 14k questions/s that is the slowest stage of the drop-in
 * path by 3x. Here the per-row value objects are created once and shared by the k lists (they are immutable ints,
 * floats, strings or None), every k still gets its own dict objects. Host code only: no CUDA, no numpy headers (the
 * arrays arrive through the buffer protocol). evi-rag_b200/writer.py falls back to its Python loop when this module
 * has not been built; tests/test_host_cpu.py checks both produce equal records. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

typedef struct {
  Py_buffer view;
  int ok;
} Buf;

static int get_buf(PyObject* obj, Buf* b, Py_ssize_t itemsize, const char* name) {
  b->ok = 0;
  if (obj == Py_None) return 1;
  if (PyObject_GetBuffer(obj, &b->view, PyBUF_C_CONTIGUOUS | PyBUF_FORMAT) != 0) return 0;
  b->ok = 1;
  if (b->view.itemsize != itemsize) {
    PyErr_Format(PyExc_TypeError, "%s: item size %zd, expected %zd", name, b->view.itemsize, itemsize);
    return 0;
  }
  return 1;
}

static PyObject* lookup_text(PyObject* map, PyObject* key) { /* new reference */
  if (map == Py_None) Py_RETURN_NONE;
  PyObject* v = PyDict_GetItemWithError(map, key); /* borrowed */
  if (v == NULL) {
    if (PyErr_Occurred()) return NULL;
    Py_RETURN_NONE;
  }
  Py_INCREF(v);
  return v;
}

/* build(count i32 [B], head i64 [B,K], relation i64 [B,K], tail i64 [B,K], score f32 [B,K], label f32 [B,K],
 *       logit_fwd f32 [B,K] | None, logit_bwd f32 [B,K] | None, K, ks (tuple of ints), entity_map | None,
 *       relation_map | None) -> list of length B: dict k -> list of edge dicts, or None when count <= 0 */
static PyObject* build(PyObject* self, PyObject* args) {
  PyObject *o_cnt, *o_h, *o_r, *o_t, *o_s, *o_l, *o_f, *o_b, *ks, *ent, *rel;
  Py_ssize_t K;
  if (!PyArg_ParseTuple(args, "OOOOOOOOnOOO", &o_cnt, &o_h, &o_r, &o_t, &o_s, &o_l, &o_f, &o_b, &K, &ks, &ent, &rel))
    return NULL;
  if (!PyTuple_Check(ks) || (ent != Py_None && !PyDict_Check(ent)) || (rel != Py_None && !PyDict_Check(rel))) {
    PyErr_SetString(PyExc_TypeError, "ks must be a tuple; entity_map / relation_map a dict or None");
    return NULL;
  }
  Buf cnt, h, r, t, s, l, f, b;
  Buf* all[8] = {&cnt, &h, &r, &t, &s, &l, &f, &b};
  for (int i = 0; i < 8; ++i) all[i]->ok = 0;
  PyObject* out = NULL;
  static PyObject* keys[11] = {0};
  static const char* names[11] = {"head_entity_id", "relation_id", "tail_entity_id", "head_text", "relation_text",
                                  "tail_text", "score", "label", "rank", "logit_fwd", "logit_bwd"};
  if (keys[0] == NULL)
    for (int i = 0; i < 11; ++i) keys[i] = PyUnicode_InternFromString(names[i]);
  if (!get_buf(o_cnt, &cnt, 4, "count") || !get_buf(o_h, &h, 8, "head_id") || !get_buf(o_r, &r, 8, "relation_id") ||
      !get_buf(o_t, &t, 8, "tail_id") || !get_buf(o_s, &s, 4, "score") || !get_buf(o_l, &l, 4, "label") ||
      !get_buf(o_f, &f, 4, "logit_fwd") || !get_buf(o_b, &b, 4, "logit_bwd"))
    goto done;
  if (!cnt.ok || !h.ok || !r.ok || !t.ok || !s.ok || !l.ok) {
    PyErr_SetString(PyExc_TypeError, "count, head_id, relation_id, tail_id, score and label are required");
    goto done;
  }
  {
    const Py_ssize_t B = cnt.view.len / 4;
    const Py_ssize_t need8 = B * K * 8, need4 = B * K * 4;
    if (K <= 0 || h.view.len < need8 || r.view.len < need8 || t.view.len < need8 || s.view.len < need4 ||
        l.view.len < need4 || (f.ok && f.view.len < need4) || (b.ok && b.view.len < need4)) {
      PyErr_SetString(PyExc_ValueError, "arrays are smaller than [B, K]");
      goto done;
    }
    const int32_t* pc = (const int32_t*)cnt.view.buf;
    const int64_t *ph = (const int64_t*)h.view.buf, *pr = (const int64_t*)r.view.buf, *pt = (const int64_t*)t.view.buf;
    const float *ps = (const float*)s.view.buf, *pl = (const float*)l.view.buf;
    const float* pf = f.ok ? (const float*)f.view.buf : NULL;
    const float* pb = b.ok ? (const float*)b.view.buf : NULL;
    const Py_ssize_t nk = PyTuple_GET_SIZE(ks);
    const int nvals = 9 + (pf != NULL) + (pb != NULL);
    out = PyList_New(B);
    if (out == NULL) goto done;
    PyObject** vals = (PyObject**)PyMem_Malloc(sizeof(PyObject*) * (size_t)K * 11);
    if (vals == NULL) {
      Py_CLEAR(out);
      PyErr_NoMemory();
      goto done;
    }
    for (Py_ssize_t q = 0; q < B; ++q) {
      Py_ssize_t kt = pc[q] < K ? pc[q] : K;
      if (kt <= 0) {
        Py_INCREF(Py_None);
        PyList_SET_ITEM(out, q, Py_None);
        continue;
      }
      int failed = 0;
      Py_ssize_t made = 0;
      for (Py_ssize_t j = 0; j < kt && !failed; ++j, ++made) { /* the row's value objects, shared by every k list */
        const Py_ssize_t i = q * K + j;
        PyObject** v = vals + j * 11;
        for (int c = 0; c < 11; ++c) v[c] = NULL;
        v[0] = PyLong_FromLongLong(ph[i]);
        v[1] = PyLong_FromLongLong(pr[i]);
        v[2] = PyLong_FromLongLong(pt[i]);
        if (v[0] && v[1] && v[2]) {
          v[3] = lookup_text(ent, v[0]);
          v[4] = lookup_text(rel, v[1]);
          v[5] = lookup_text(ent, v[2]);
        }
        v[6] = PyFloat_FromDouble((double)ps[i]);
        v[7] = PyFloat_FromDouble((double)pl[i]);
        v[8] = PyLong_FromSsize_t(j + 1);
        int c = 9;
        if (pf) v[c++] = PyFloat_FromDouble((double)pf[i]);
        if (pb) v[c++] = PyFloat_FromDouble((double)pb[i]);
        for (int x = 0; x < nvals; ++x)
          if (v[x] == NULL) failed = 1;
      }
      PyObject* by_k = failed ? NULL : PyDict_New();
      for (Py_ssize_t a = 0; by_k != NULL && a < nk; ++a) {
        PyObject* kobj = PyTuple_GET_ITEM(ks, a);
        const Py_ssize_t kv = PyLong_AsSsize_t(kobj);
        if (kv < 0 && PyErr_Occurred()) {
          Py_CLEAR(by_k);
          break;
        }
        const Py_ssize_t n = kv < kt ? kv : kt;
        PyObject* lst = PyList_New(n);
        if (lst == NULL) {
          Py_CLEAR(by_k);
          break;
        }
        for (Py_ssize_t j = 0; j < n; ++j) {
          PyObject* d = PyDict_New();
          PyObject** v = vals + j * 11;
          int bad = d == NULL;
          for (int c = 0; !bad && c < 9; ++c) bad = PyDict_SetItem(d, keys[c], v[c]) != 0;
          int c = 9;
          if (!bad && pf) bad = PyDict_SetItem(d, keys[9], v[c++]) != 0;
          if (!bad && pb) bad = PyDict_SetItem(d, keys[10], v[c++]) != 0;
          if (bad) {
            Py_XDECREF(d);
            Py_CLEAR(lst);
            break;
          }
          PyList_SET_ITEM(lst, j, d);
        }
        if (lst == NULL || PyDict_SetItem(by_k, kobj, lst) != 0) {
          Py_XDECREF(lst);
          Py_CLEAR(by_k);
          break;
        }
        Py_DECREF(lst);
      }
      for (Py_ssize_t j = 0; j < made; ++j)
        for (int c = 0; c < 11; ++c) Py_XDECREF(vals[j * 11 + c]);
      if (by_k == NULL) {
        if (!PyErr_Occurred()) PyErr_NoMemory();
        Py_CLEAR(out);
        break;
      }
      PyList_SET_ITEM(out, q, by_k);
    }
    PyMem_Free(vals);
  }
done:
  for (int i = 0; i < 8; ++i)
    if (all[i]->ok) PyBuffer_Release(&all[i]->view);
  return out;
}

static PyMethodDef methods[] = {
    {"build", build, METH_VARARGS, "triplets_by_k dictionaries of one batch from its [B, K] host columns"},
    {NULL, NULL, 0, NULL}};
static struct PyModuleDef module = {PyModuleDef_HEAD_INIT, "_evi_records", NULL, -1, methods};
PyMODINIT_FUNC PyInit__evi_records(void) { return PyModule_Create(&module); }
