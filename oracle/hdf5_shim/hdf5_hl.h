// TEST INFRASTRUCTURE ONLY (oracle build). "High-level" (H5LT) part of the
// in-memory HDF5 stand-in; see hdf5.h in this directory for what and why.
#ifndef SGX_ORACLE_HDF5_HL_SHIM_H_
#define SGX_ORACLE_HDF5_HL_SHIM_H_

#include "hdf5.h"

namespace h5shim {
inline const Attr *find_attr(hid_t loc, const char *obj, const char *name) {
  Node *n = resolve(from_id(loc), obj);
  if (!n) return nullptr;
  auto it = n->attrs.find(name);
  return it == n->attrs.end() ? nullptr : &it->second;
}
} // namespace h5shim

inline herr_t H5LTget_attribute_int(hid_t loc, const char *obj, const char *name,
                                    int *out) {
  const h5shim::Attr *a = h5shim::find_attr(loc, obj, name);
  if (!a || a->cls != H5T_INTEGER) return -1;
  for (std::size_t k = 0; k < a->i.size(); ++k)
    out[k] = a->i[k];
  return 0;
}
inline herr_t H5LTget_attribute_double(hid_t loc, const char *obj, const char *name,
                                       double *out) {
  const h5shim::Attr *a = h5shim::find_attr(loc, obj, name);
  if (!a || a->cls != H5T_FLOAT) return -1;
  for (std::size_t k = 0; k < a->d.size(); ++k)
    out[k] = a->d[k];
  return 0;
}
inline herr_t H5LTget_attribute_char(hid_t loc, const char *obj, const char *name,
                                     char *out) {
  const h5shim::Attr *a = h5shim::find_attr(loc, obj, name);
  if (!a || a->cls != H5T_STRING) return -1;
  std::memcpy(out, a->s.data(), a->s.size());
  return 0;
}
inline herr_t H5LTget_attribute_string(hid_t loc, const char *obj, const char *name,
                                       char *out) {
  const h5shim::Attr *a = h5shim::find_attr(loc, obj, name);
  if (!a || a->cls != H5T_STRING) return -1;
  std::memcpy(out, a->s.c_str(), a->s.size() + 1);
  return 0;
}
inline herr_t H5LTget_attribute_info(hid_t loc, const char *obj, const char *name,
                                     hsize_t *dims, H5T_class_t *cls,
                                     std::size_t *type_size) {
  const h5shim::Attr *a = h5shim::find_attr(loc, obj, name);
  if (!a) return -1;
  *cls = a->cls;
  if (a->cls == H5T_STRING) {
    dims[0] = 1;
    *type_size = a->s.size();
  } else if (a->cls == H5T_INTEGER) {
    dims[0] = a->i.size();
    *type_size = sizeof(int);
  } else {
    dims[0] = a->d.size();
    *type_size = sizeof(double);
  }
  return 0;
}
inline herr_t H5LTread_dataset(hid_t loc, const char *name, hid_t type, void *buf) {
  h5shim::Node *n = h5shim::resolve(h5shim::from_id(loc), name);
  if (!n || !n->is_dataset) return -1;
  if (type == H5T_NATIVE_DOUBLE) {
    std::memcpy(buf, n->data.data(), n->data.size() * sizeof(double));
  } else if (type == H5T_NATIVE_FLOAT) {
    float *f = static_cast<float *>(buf);
    for (std::size_t k = 0; k < n->data.size(); ++k)
      f[k] = static_cast<float>(n->data[k]);
  } else {
    return -1;
  }
  return 0;
}
inline herr_t H5LTmake_dataset(hid_t loc, const char *name, int rank, const hsize_t *dims,
                               hid_t type, const void *buf) {
  h5shim::Node *p = h5shim::from_id(loc);
  if (!p) return -1;
  std::size_t n = 1;
  for (int r = 0; r < rank; ++r)
    n *= dims[r];
  if (type == H5T_NATIVE_DOUBLE) {
    h5shim::set_dataset(p, name, static_cast<const double *>(buf), n);
  } else if (type == H5T_NATIVE_FLOAT) {
    std::vector<double> tmp(n);
    const float *f = static_cast<const float *>(buf);
    for (std::size_t k = 0; k < n; ++k)
      tmp[k] = f[k];
    h5shim::set_dataset(p, name, tmp.data(), n);
  } else {
    return -1;
  }
  return 0;
}
inline herr_t H5LTset_attribute_int(hid_t loc, const char *obj, const char *name,
                                    const int *buf, std::size_t n) {
  h5shim::Node *nd = h5shim::resolve(h5shim::from_id(loc), obj);
  if (!nd) return -1;
  h5shim::set_attr_int(nd, name, std::vector<int>(buf, buf + n));
  return 0;
}
inline herr_t H5LTset_attribute_double(hid_t loc, const char *obj, const char *name,
                                       const double *buf, std::size_t n) {
  h5shim::Node *nd = h5shim::resolve(h5shim::from_id(loc), obj);
  if (!nd) return -1;
  h5shim::Attr a;
  a.cls = H5T_FLOAT;
  a.d.assign(buf, buf + n);
  nd->attrs[name] = a;
  return 0;
}
inline herr_t H5LTset_attribute_string(hid_t loc, const char *obj, const char *name,
                                       const char *val) {
  h5shim::Node *nd = h5shim::resolve(h5shim::from_id(loc), obj);
  if (!nd) return -1;
  h5shim::Attr a;
  a.cls = H5T_STRING;
  a.s = val;
  nd->attrs[name] = a;
  return 0;
}

#endif // SGX_ORACLE_HDF5_HL_SHIM_H_
