// TEST INFRASTRUCTURE ONLY (oracle). Never linked into or called by the product path.
//
// Second translation unit of the `annonet_ref` oracle driver: it instantiates the
// reference's own `net_type` (dlib-dnn-pimpl-wrapper/NetStructure.h:172-234) from the
// sources under /root/reference so that quantities the pimpl API hides can be dumped:
// pristine random-init parameters, the training-mode logits, the loss value, the
// per-layer parameter gradients (dlib/dlib/dnn/core.h:3806-3816) and, for debugging,
// every layer's output tensor.  Nothing here re-implements any arithmetic: every number
// comes out of dlib's CPU code.
#include "dlib-dnn-pimpl-wrapper/NetPimpl.h"
#include "dlib-dnn-pimpl-wrapper/NetStructure.h"
#include "cli_io.h"

#include <cstdlib>
#include <fstream>
#include <iostream>

namespace {

struct output_dumper {
    std::ostream& out;
    explicit output_dumper(std::ostream& o) : out(o) {}

    template <typename T, typename U, typename E>
    void operator()(size_t idx, dlib::add_layer<T, U, E>& l) const
    {
        // layers under an in-place layer (affine_) refuse get_output() (core.h:1102-1109): skip them
        try { ref_io::write_tensor(out, static_cast<int>(idx), l.get_output()); } catch (const std::exception&) {}
    }

    template <typename T>
    void operator()(size_t, T&) const {}
};

}  // namespace

// Materialise the parameters of a configured-but-unallocated net_type stream (as written by
// TrainingNet::Serialize right after SetNetWidth/SetClassCount) exactly the way dlib does on
// the first forward pass (layers.h:170-193 `con_::setup`: Glorot-uniform from
// dlib::rand(std::rand())), with batch size 1 so that no bn_ running statistic moves
// (layers.h:1392).
int ref_materialize(const std::string& in_path, const std::string& out_path, unsigned seed)
{
    net_type net;
    {
        std::ifstream in(in_path, std::ios::binary);
        dlib::deserialize(net, in);
    }
    std::srand(seed);
    NetPimpl::input_type img(127, 127);
    for (auto& p : img) { p.red = 128; p.green = 128; p.blue = 128; }
    dlib::resizable_tensor x;
    net.to_tensor(&img, &img + 1, x);
    net.subnet().forward(x);
    net.clean();
    std::ofstream out(out_path, std::ios::binary);
    dlib::serialize(net, out);
    return 0;
}

// loss + gradients of ONE minibatch on the given net_type stream, without a parameter update:
// add_loss_layer::compute_parameter_gradients (core.h:2570-2580).
int ref_grads(const std::string& net_path, const std::vector<NetPimpl::input_type>& images,
              const std::vector<NetPimpl::training_label_type>& labels, const std::string& out_prefix, bool dump_layers)
{
    net_type net;
    {
        std::ifstream in(net_path, std::ios::binary);
        dlib::deserialize(net, in);
    }
    dlib::resizable_tensor x;
    net.to_tensor(images.begin(), images.end(), x);
    const double loss = net.compute_parameter_gradients(x, labels.begin());

    {
        std::ofstream out(out_prefix + ".loss", std::ios::binary);
        out.write(reinterpret_cast<const char*>(&loss), sizeof(loss));
    }
    {
        std::ofstream out(out_prefix + ".logits", std::ios::binary);
        ref_io::write_tensor(out, 1, net.subnet().get_output());
    }
    {
        std::ofstream out(out_prefix + ".grads", std::ios::binary);
        dlib::visit_layer_parameter_gradients(net, [&](size_t idx, dlib::tensor& t) { ref_io::write_tensor(out, static_cast<int>(idx), t); });
    }
    if (dump_layers) {
        std::ofstream out(out_prefix + ".layers", std::ios::binary);
        dlib::visit_layers(net, output_dumper(out));
    }
    std::cout << "loss " << loss << std::endl;
    return 0;
}

// Same, inference net (anet_type): dump every layer output of one forward pass.
int ref_forward_layers(const std::string& anet_path, const NetPimpl::input_type& image, const std::string& out_path)
{
    anet_type anet;
    {
        std::ifstream in(anet_path, std::ios::binary);
        dlib::deserialize(anet, in);
    }
    dlib::resizable_tensor x;
    anet.to_tensor(&image, &image + 1, x);
    anet.subnet().forward(x);
    std::ofstream out(out_path, std::ios::binary);
    dlib::visit_layers(anet, output_dumper(out));
    return 0;
}
