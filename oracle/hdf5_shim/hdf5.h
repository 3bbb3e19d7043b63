// TEST INFRASTRUCTURE ONLY (oracle build). Not part of the product path.
//
// In-memory stand-in for the handful of HDF5 C calls the reference's Spiner
// EOS constructors make ("Plan A").  libhdf5 is not installed in
// this image, and SpinerEOSDependsRhoT / SpinerEOSDependsRhoSie can only be
// constructed from an .sp5 (HDF5) file:
//   singularity-eos/eos/eos_spiner_rho_temp.hpp:373-424, :545-687
//   singularity-eos/eos/eos_spiner_rho_sie.hpp:434-577
//   utils/spiner/spiner/databox.hpp:853-906 (DataBox::loadHDF)
//   utils/spiner/spiner/piecewise_grid_1d.hpp:181-207, regular_grid_1d.hpp:137-154
// The harness (oracle/ref_harness.cpp) fills a tree of groups / attributes /
// datasets through the h5shim:: builder below, registers it under a file name,
// and the *unmodified* reference constructors then read it through these calls.
#ifndef SGX_ORACLE_HDF5_SHIM_H_
#define SGX_ORACLE_HDF5_SHIM_H_

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

typedef std::int64_t hid_t;
typedef int herr_t;
typedef unsigned long long hsize_t;
typedef int htri_t;

#define H5P_DEFAULT 0
#define H5E_DEFAULT 0
#define H5F_ACC_RDONLY 0u
#define H5F_ACC_TRUNC 2u
#define H5T_NATIVE_DOUBLE 1
#define H5T_NATIVE_FLOAT 2
#define H5T_NATIVE_INT 3

typedef enum H5T_class_t {
  H5T_NO_CLASS = -1,
  H5T_INTEGER = 0,
  H5T_FLOAT = 1,
  H5T_STRING = 3
} H5T_class_t;

typedef herr_t (*H5E_auto2_t)(hid_t estack, void *client_data);

namespace h5shim {

struct Attr {
  H5T_class_t cls = H5T_NO_CLASS;
  std::vector<int> i;
  std::vector<double> d;
  std::string s;
};

struct Node {
  bool is_dataset = false;
  std::map<std::string, std::unique_ptr<Node>> children;
  std::map<std::string, Attr> attrs;
  std::vector<double> data; // datasets are always double here
  Node *child(const std::string &name, bool create) {
    auto it = children.find(name);
    if (it != children.end()) return it->second.get();
    if (!create) return nullptr;
    children[name] = std::make_unique<Node>();
    return children[name].get();
  }
};

inline std::map<std::string, std::unique_ptr<Node>> &files() {
  static std::map<std::string, std::unique_ptr<Node>> f;
  return f;
}
inline std::vector<Node *> &handles() {
  static std::vector<Node *> h(1, nullptr); // id 0 unused
  return h;
}
inline hid_t to_id(Node *n) {
  handles().push_back(n);
  return static_cast<hid_t>(handles().size() - 1);
}
inline Node *from_id(hid_t id) {
  if (id <= 0 || static_cast<std::size_t>(id) >= handles().size()) return nullptr;
  return handles()[id];
}
// Resolve "a/b/c", "." and "/" relative to loc.
inline Node *resolve(Node *loc, const char *path, bool create = false) {
  if (loc == nullptr) return nullptr;
  std::string p(path);
  Node *cur = loc;
  std::size_t pos = 0;
  while (pos <= p.size()) {
    std::size_t nxt = p.find('/', pos);
    if (nxt == std::string::npos) nxt = p.size();
    std::string tok = p.substr(pos, nxt - pos);
    pos = nxt + 1;
    if (tok.empty() || tok == ".") continue;
    cur = cur->child(tok, create);
    if (cur == nullptr) return nullptr;
  }
  return cur;
}
inline Node *file_root(const std::string &name, bool create) {
  auto it = files().find(name);
  if (it != files().end()) return it->second.get();
  if (!create) return nullptr;
  files()[name] = std::make_unique<Node>();
  return files()[name].get();
}
inline void drop_file(const std::string &name) { files().erase(name); }

// ---- builder helpers used by the harness -------------------------------
inline void set_attr_int(Node *n, const std::string &name, const std::vector<int> &v) {
  Attr a;
  a.cls = H5T_INTEGER;
  a.i = v;
  n->attrs[name] = a;
}
inline void set_attr_double(Node *n, const std::string &name, double v) {
  Attr a;
  a.cls = H5T_FLOAT;
  a.d = {v};
  n->attrs[name] = a;
}
inline void set_dataset(Node *parent, const std::string &name, const double *v,
                        std::size_t n) {
  Node *d = parent->child(name, true);
  d->is_dataset = true;
  d->data.assign(v, v + n);
}
} // namespace h5shim

// ---- the C API subset ----------------------------------------------------
inline herr_t H5Eset_auto(hid_t, H5E_auto2_t, void *) { return 0; }
inline herr_t H5Eprint2(hid_t, FILE *) { return 0; }
#define H5Eprint H5Eprint2

inline hid_t H5Fopen(const char *name, unsigned, hid_t) {
  h5shim::Node *r = h5shim::file_root(name, false);
  return r ? h5shim::to_id(r) : -1;
}
inline hid_t H5Fcreate(const char *name, unsigned, hid_t, hid_t) {
  h5shim::drop_file(name);
  return h5shim::to_id(h5shim::file_root(name, true));
}
inline herr_t H5Fclose(hid_t) { return 0; }
inline hid_t H5Gopen(hid_t loc, const char *name, hid_t) {
  h5shim::Node *n = h5shim::resolve(h5shim::from_id(loc), name);
  return n ? h5shim::to_id(n) : -1;
}
inline hid_t H5Gcreate(hid_t loc, const char *name, hid_t, hid_t, hid_t) {
  h5shim::Node *n = h5shim::resolve(h5shim::from_id(loc), name, true);
  return n ? h5shim::to_id(n) : -1;
}
inline herr_t H5Gclose(hid_t) { return 0; }
inline htri_t H5Lexists(hid_t loc, const char *name, hid_t) {
  return h5shim::resolve(h5shim::from_id(loc), name) != nullptr ? 1 : 0;
}
inline htri_t H5Aexists_by_name(hid_t loc, const char *obj, const char *attr, hid_t) {
  h5shim::Node *n = h5shim::resolve(h5shim::from_id(loc), obj);
  if (!n) return 0;
  return n->attrs.count(attr) ? 1 : 0;
}

#endif // SGX_ORACLE_HDF5_SHIM_H_
