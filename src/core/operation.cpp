#include "tenzing/operation.hpp"

#include <mutex>
#include <unordered_map>

#include "tenzing/cuda/ops_cuda.hpp"

namespace tenzing {

namespace {
struct KeyTable {
  std::mutex mu;
  std::unordered_map<std::string, uint32_t> ids;
  std::vector<const std::string *> keys{nullptr}; // uid 0 is reserved for "not interned"
};
KeyTable &table() {
  static KeyTable *t = new KeyTable(); // leaked on purpose: ops may outlive static dtors
  return *t;
}
} // namespace

uint32_t intern_key(const std::string &key) {
  KeyTable &t = table();
  std::lock_guard<std::mutex> lk(t.mu);
  auto it = t.ids.find(key);
  if (it != t.ids.end()) return it->second;
  const uint32_t id = static_cast<uint32_t>(t.keys.size());
  auto ins = t.ids.emplace(key, id);
  t.keys.push_back(&ins.first->first);
  return id;
}

const std::string &interned_key(uint32_t uid) {
  KeyTable &t = table();
  std::lock_guard<std::mutex> lk(t.mu);
  if (uid == 0 || uid >= t.keys.size()) THROW_RUNTIME("unknown op uid " << uid);
  return *t.keys[uid];
}

Json OpBase::json() const {
  Json j = Json::object();
  j["name"] = name();
  return j;
}

Json NoOp::json() const {
  Json j = Json::object();
  j["name"] = name();
  j["kind"] = "NoOp";
  return j;
}

void keep_uniques(std::vector<BoundOpPtr> &v) {
  std::vector<BoundOpPtr> out;
  for (const auto &e : v) {
    bool dup = false;
    for (const auto &o : out) {
      if (o->uid() == e->uid()) {
        dup = true;
        break;
      }
    }
    if (!dup) out.push_back(e);
  }
  v.swap(out);
}

std::vector<BoundOpPtr> make_platform_variations(const Platform &plat, const OpPtr &op) {
  std::vector<BoundOpPtr> r;
  if (auto g = std::dynamic_pointer_cast<GpuOp>(op)) {
    for (const Stream &s : plat.streams_) r.push_back(std::make_shared<BoundGpuOp>(g, s));
  } else if (auto b = std::dynamic_pointer_cast<BoundOp>(op)) {
    r.push_back(b);
  } else {
    THROW_RUNTIME("make_platform_variations: " << op->desc() << " is neither GpuOp nor BoundOp");
  }
  return r;
}

bool unbound_contains(const std::vector<BoundOpPtr> &v, const OpPtr &e) {
  for (const auto &x : v)
    if (x->unbound_uid() == e->unbound_uid()) return true;
  return false;
}

bool contains(const std::vector<OpPtr> &v, const OpPtr &e) {
  for (const auto &x : v)
    if (x->uid() == e->uid()) return true;
  return false;
}

} // namespace tenzing
