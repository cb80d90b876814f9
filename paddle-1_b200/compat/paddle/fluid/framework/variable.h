// compat: framework/variable.h — type-erased holder with IsType/Get/GetMutable (variable.h:27-92).
#pragma once
#include <memory>
#include <typeindex>
#include "paddle/fluid/platform/enforce.h"

namespace paddle {
namespace framework {

class Variable {
 public:
  template <typename T>
  const T& Get() const {
    PADDLE_ENFORCE(holder_ != nullptr, "Variable must hold some thing");
    PADDLE_ENFORCE(IsType<T>(), "Variable must be type %s, the holding type is %s", typeid(T).name(),
                   holder_->type.name());
    return *static_cast<const T*>(holder_->ptr());
  }
  bool IsInitialized() const { return holder_ != nullptr; }
  template <typename T>
  T* GetMutable() {
    if (!IsType<T>()) holder_.reset(new HolderImpl<T>());
    return static_cast<T*>(holder_->ptr());
  }
  template <typename T>
  bool IsType() const { return holder_ != nullptr && holder_->type == std::type_index(typeid(T)); }
  void Clear() { holder_.reset(); }
  std::type_index Type() const {
    PADDLE_ENFORCE(holder_ != nullptr, "Must hold memory");
    return holder_->type;
  }

 private:
  struct Holder {
    explicit Holder(std::type_index t) : type(t) {}
    virtual ~Holder() {}
    virtual void* ptr() = 0;
    std::type_index type;
  };
  template <typename T>
  struct HolderImpl : Holder {
    HolderImpl() : Holder(std::type_index(typeid(T))) {}
    void* ptr() override { return &obj; }
    T obj;
  };
  std::unique_ptr<Holder> holder_;
};

}  // namespace framework
}  // namespace paddle
