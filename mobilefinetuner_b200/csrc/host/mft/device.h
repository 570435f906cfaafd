// Device tag of ops::Tensor (API mirror of the reference's core/device.h:8-98). In this implementation the
// tag is only a label: every tensor's storage is HBM-resident with a lazily synchronised host mirror, so
// tensors created as kCPU by the unchanged trainers still compute on the B200.
#pragma once
#include <cstdint>
#include <string>

namespace ops {

enum class DeviceType : int8_t { kCPU = 0, kCUDA = 1, kMetal = 2 };

class Device {
public:
    Device(DeviceType type, int index = 0) : type_(type), index_(index) {}
    DeviceType type() const { return type_; }
    int index() const { return index_; }
    bool is_cpu() const { return type_ == DeviceType::kCPU; }
    bool is_cuda() const { return type_ == DeviceType::kCUDA; }
    bool is_metal() const { return type_ == DeviceType::kMetal; }
    std::string to_string() const {
        if (is_cpu()) return "cpu";
        return std::string(is_cuda() ? "cuda:" : "metal:") + std::to_string(index_);
    }
    bool operator==(const Device& o) const { return type_ == o.type_ && index_ == o.index_; }
    bool operator!=(const Device& o) const { return !(*this == o); }

private:
    DeviceType type_;
    int index_;
};

struct DeviceManager {
    static Device best_available() { return Device(DeviceType::kCUDA, 0); }
    static Device cpu() { return Device(DeviceType::kCPU, 0); }
    static Device cuda(int index = 0) { return Device(DeviceType::kCUDA, index); }
    static Device metal(int index = 0) { return Device(DeviceType::kMetal, index); }
    static bool cuda_available() { return true; }
    static bool metal_available() { return false; }
    static int cuda_device_count() { return 1; }
    static int metal_device_count() { return 0; }
};

static const Device kCPU(DeviceType::kCPU, 0);
static const Device kCUDA(DeviceType::kCUDA, 0);
static const Device kMetal(DeviceType::kMetal, 0);

}  // namespace ops
