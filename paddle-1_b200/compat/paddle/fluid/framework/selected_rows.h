// compat: framework/selected_rows.h — {rows_: Vector<int64_t> (duplicates allowed), value_: Tensor
// [rows.size(), D...], height_}.  Mirrors selected_rows.h:32-140 of the reference; the auto-grown
// sparse-table API is limited to Index()/AutoGrownIndex() as the sgd SelectedRows-param branch uses.
#pragma once
#include <algorithm>
#include <memory>
#include <vector>
#include "paddle/fluid/framework/lod_tensor.h"
#include "paddle/fluid/framework/mixed_vector.h"

namespace paddle {
namespace framework {

class SelectedRows {
 public:
  SelectedRows(const std::vector<int64_t>& rows, const int64_t& height)
      : rows_(rows), height_(height) { value_.reset(new Tensor()); }
  SelectedRows() : height_(0) { value_.reset(new Tensor()); }

  platform::Place place() const { return value_->place(); }
  const Tensor& value() const { return *value_; }
  Tensor* mutable_value() { return value_.get(); }
  int64_t height() const { return height_; }
  void set_height(int64_t height) { height_ = height; }
  const Vector<int64_t>& rows() const { return rows_; }
  Vector<int64_t>* mutable_rows() { return &rows_; }
  void set_rows(const Vector<int64_t>& rows) { rows_ = rows; }

  // first position of `key` in rows, -1 when absent (selected_rows.h:83-89)
  int64_t Index(int64_t key) const {
    auto it = std::find(rows_.begin(), rows_.end(), key);
    if (it == rows_.end()) return static_cast<int64_t>(-1);
    return static_cast<int64_t>(std::distance(rows_.begin(), it));
  }
  bool HasKey(int64_t key) const { return Index(key) >= 0; }
  int64_t AutoGrownIndex(int64_t key, bool auto_grown) {
    int64_t idx = Index(key);
    if (idx >= 0) return idx;
    PADDLE_ENFORCE(auto_grown, "key %ld not found", static_cast<long>(key));
    PADDLE_ENFORCE_LT(static_cast<int64_t>(rows_.size()), value_->dims()[0], "selected rows is full");
    rows_.push_back(key);
    return static_cast<int64_t>(rows_.size()) - 1;
  }
  DDim GetCompleteDims() const {
    std::vector<int64_t> dims = vectorize(value_->dims());
    dims[0] = height_;
    return make_ddim(dims);
  }

 private:
  Vector<int64_t> rows_;
  std::unique_ptr<Tensor> value_{nullptr};
  int64_t height_;
};

}  // namespace framework
}  // namespace paddle
