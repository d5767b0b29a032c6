// bear_b200._api -- pybind11 module shaped like the reference's `visr_bear.api`
// (visr_bear/pythonwrappers/api.cpp:23-181) over include/bear_b200/host.hpp, i.e. over the CUDA render
// core.  Same class and member names, same `process` buffer checks (array_conversion.hpp:12-44:
// float32, 2-D, wrong length / wrong number of channels / sample stride -> ValueError), same return
// values of add_*_block (False = too early, push again after the next process()).
//
// One difference, by necessity: ADM metadata -> gains is libear's job (ear::GainCalculatorObjects etc.,
// which need Boost and are not part of this repository), so ObjectsInput and DirectSpeakersInput carry
// the GAIN VECTORS libear would produce (`direct_gains` / `diffuse_gains`, `gains`; length = number of
// virtual loudspeakers) instead of `type_metadata`.  HOAInput is complete: channels + type_metadata
// {orders, degrees, normalization}.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <optional>
#include <string>
#include <vector>

#include "../../include/bear_b200/host.hpp"

namespace py = pybind11;
using namespace bear_b200;

namespace {

// array_conversion.hpp:12-44
std::vector<float *> array_to_pointers(py::array_t<float, 0> &arr, size_t n_channels, size_t n_samples,
                                       const std::string &name)
{
  if (arr.ndim() != 2) throw py::value_error(name + " must have 2 dimensions");
  if ((size_t)arr.shape(1) != n_samples) throw py::value_error(name + " has wrong length");
  const bool stride_invalid = arr.shape(0) == 0 || arr.shape(1) == 1;
  if (!stride_invalid && (size_t)arr.strides(1) != sizeof(float))
    throw py::value_error(name + " must have stride of 1 along samples");
  std::vector<float *> ptrs((size_t)arr.shape(0));
  for (size_t i = 0; i < ptrs.size(); i++) ptrs[i] = arr.mutable_data(i, 0);
  if (ptrs.size() != n_channels) throw py::value_error(name + " has wrong number of channels");
  return ptrs;
}

struct HOATypeMetadata {
  std::vector<int> orders, degrees;
  std::string normalization = "SN3D";
};
// the reference's HOAInput: channels + type_metadata (api.hpp:108-112)
struct PyHOAInput : MetadataInput {
  std::vector<size_t> channels;
  HOATypeMetadata type_metadata;
};

struct DataFileMetadata {
  bool has_metadata = false, is_released = false;
  std::string label, description;
  static DataFileMetadata read_from_file(const std::string &path)   // api.cpp:330-353
  {
    bear_cuda_data_file_info info;
    if (bear_cuda_read_data_file_info(path.c_str(), &info) != BEAR_CUDA_OK)
      throw std::runtime_error(bear_cuda_last_error(nullptr));
    DataFileMetadata m;
    m.has_metadata = info.has_metadata != 0;
    m.is_released = info.is_released != 0;
    m.label = info.label;
    m.description = info.description;
    return m;
  }
};

struct RendererWrapper {
  explicit RendererWrapper(const Config &c) : config(c), renderer(c) {}
  Config config;
  Renderer renderer;
};

}  // namespace

PYBIND11_MODULE(_api, m)
{
  m.doc() = "visr_bear.api-shaped binding of the CUDA BEAR render core (gain vectors instead of ADM type metadata)";

  py::class_<Config>(m, "Config")
      .def(py::init<>())
      .def(py::init<Config>())
      .def_property("num_objects_channels", &Config::get_num_objects_channels, &Config::set_num_objects_channels)
      .def_property("num_direct_speakers_channels", &Config::get_num_direct_speakers_channels,
                    &Config::set_num_direct_speakers_channels)
      .def_property("num_hoa_channels", &Config::get_num_hoa_channels, &Config::set_num_hoa_channels)
      .def_property("period_size", &Config::get_period_size, &Config::set_period_size)
      .def_property("sample_rate", &Config::get_sample_rate, &Config::set_sample_rate)
      .def_property("data_path", &Config::get_data_path, &Config::set_data_path)
      .def_property("fft_implementation", &Config::get_fft_implementation, &Config::set_fft_implementation)
      .def_property("device", &Config::get_device, &Config::set_device)   // CUDA device ordinal (no counterpart)
      .def("validate", &Config::validate);

  py::class_<Time>(m, "Time")
      .def(py::init<int64_t>())
      .def(py::init<int64_t, int64_t>())
      .def_property_readonly("numerator", &Time::numerator)
      .def_property_readonly("denominator", &Time::denominator)
      .def("__eq__", [](const Time &a, const Time &b) { return a == b; })
      .def("__repr__", [](const Time &t) {
        return "Time(" + std::to_string(t.numerator()) + ", " + std::to_string(t.denominator()) + ")";
      });

  py::class_<MetadataInput>(m, "MetadataInput")
      .def(py::init<>())
      .def_readwrite("rtime", &MetadataInput::rtime)
      .def_readwrite("duration", &MetadataInput::duration);

  py::class_<ObjectsGains, MetadataInput>(m, "ObjectsInput")
      .def(py::init<>())
      .def(py::init<const ObjectsGains &>())
      .def_readwrite("direct_gains", &ObjectsGains::direct)
      .def_readwrite("diffuse_gains", &ObjectsGains::diffuse)
      .def_readwrite("interpolationLength", &ObjectsGains::interpolationLength);

  py::class_<DirectSpeakersGains, MetadataInput>(m, "DirectSpeakersInput")
      .def(py::init<>())
      .def(py::init<const DirectSpeakersGains &>())
      .def_readwrite("gains", &DirectSpeakersGains::gains);

  py::class_<HOATypeMetadata>(m, "HOATypeMetadata")
      .def(py::init<>())
      .def_readwrite("orders", &HOATypeMetadata::orders)
      .def_readwrite("degrees", &HOATypeMetadata::degrees)
      .def_readwrite("normalization", &HOATypeMetadata::normalization);

  py::class_<PyHOAInput, MetadataInput>(m, "HOAInput")
      .def(py::init<>())
      .def(py::init<const PyHOAInput &>())
      .def_readwrite("type_metadata", &PyHOAInput::type_metadata)
      .def_readwrite("channels", &PyHOAInput::channels);

  py::class_<Listener>(m, "Listener")
      .def(py::init<>())
      .def_property("position_cart", &Listener::get_position_cart, &Listener::set_position_cart)
      .def_property("orientation_quaternion", &Listener::get_orientation_quaternion, &Listener::set_orientation_quaternion)
      .def_property_readonly("look", &Listener::look)
      .def_property_readonly("right", &Listener::right)
      .def_property_readonly("up", &Listener::up);

  py::class_<RendererWrapper>(m, "Renderer")
      .def(py::init<const Config &>())
      .def("add_objects_block",
           [](RendererWrapper &r, size_t channel, const ObjectsGains &b) { return r.renderer.add_objects_block(channel, b); })
      .def("add_direct_speakers_block",
           [](RendererWrapper &r, size_t channel, const DirectSpeakersGains &b) {
             return r.renderer.add_direct_speakers_block(channel, b);
           })
      .def("add_hoa_block",
           [](RendererWrapper &r, size_t stream, const PyHOAInput &b) {
             HOAInput h;
             h.rtime = b.rtime; h.duration = b.duration;
             h.channels = b.channels;
             h.orders = b.type_metadata.orders; h.degrees = b.type_metadata.degrees;
             h.normalization = b.type_metadata.normalization;
             return r.renderer.add_hoa_block(stream, h);
           })
      .def(
          "process",
          [](RendererWrapper &r, py::array_t<float, 0> objects_input, py::array_t<float, 0> direct_speakers_input,
             py::array_t<float, 0> hoa_input, py::array_t<float, 0> output) {
            const size_t B = r.config.get_period_size();
            auto o = array_to_pointers(objects_input, r.config.get_num_objects_channels(), B, "objects_input");
            auto d = array_to_pointers(direct_speakers_input, r.config.get_num_direct_speakers_channels(), B,
                                       "direct_speakers_input");
            auto h = array_to_pointers(hoa_input, r.config.get_num_hoa_channels(), B, "hoa_input");
            auto out = array_to_pointers(output, 2, B, "output");
            r.renderer.process(o.data(), d.data(), h.data(), out.data());
          },
          py::arg("objects_input").noconvert(true), py::arg("direct_speakers_input").noconvert(true),
          py::arg("hoa_input").noconvert(true), py::arg("output").noconvert(true))
      .def("get_block_start_time", [](RendererWrapper &r) { return r.renderer.get_block_start_time(); })
      .def("set_block_start_time", [](RendererWrapper &r, const Time &t) { r.renderer.set_block_start_time(t); })
      .def("set_listener", [](RendererWrapper &r, const Listener &l, const std::optional<Time> &t) { r.renderer.set_listener(l, t); },
           py::arg("listener"), py::arg("interpolation_time") = std::nullopt)
      .def_property_readonly("num_gains", [](RendererWrapper &r) { return r.renderer.num_gains(); })
      .def_property_readonly("selected_view", [](RendererWrapper &r) { return r.renderer.selected_view(); })
      .def_property_readonly("residual_listener", [](RendererWrapper &r) { return r.renderer.residual_listener(); });

  py::class_<DataFileMetadata>(m, "DataFileMetadata")
      .def_static("read_from_file", &DataFileMetadata::read_from_file)
      .def_readonly("has_metadata", &DataFileMetadata::has_metadata)
      .def_readonly("label", &DataFileMetadata::label)
      .def_readonly("description", &DataFileMetadata::description)
      .def_readonly("is_released", &DataFileMetadata::is_released);
}
