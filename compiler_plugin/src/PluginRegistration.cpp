// iree_b200 compiler plugin: installs the B200 match-and-replace spec into the
// preprocessing pipeline, so matching linalg.matmul / linalg.batch_matmul /
// iree_linalg_ext.attention ops become hal.dispatch.extern ops bound to the sm_100a cubins.
//
// Shape of a plugin: samples/compiler_plugins/example/src/PluginRegistration.cpp:17-45
// (PluginSession<Derived, Options>, iree_register_compiler_plugin_<id>). The hook is
// PipelineExtensions::extendPreprocessingPassPipeline (PluginAPI/Client.h:83), invoked from
// Preprocessing/Passes.cpp:107-110 after the command-line recipes; the pass added is the
// same transform interpreter the stock --iree-preprocessing-transform-spec-filename flag adds
// (Preprocessing/Passes.cpp:78-84), pointed at the spec gen_spec.py generated.
//
// Plugins are linked statically into iree-compile, which needs the LLVM tree; this file is
// therefore not compiled in the CPU-only build container (see compiler_plugin/README.md).

#include "iree/compiler/PluginAPI/Client.h"
#include "iree/compiler/Preprocessing/Common/Passes.h"
#include "mlir/Pass/PassManager.h"

using namespace mlir;
using namespace mlir::iree_compiler;

namespace {

struct B200Options {
  // Path of the generated spec (python -m compiler_plugin.gen_spec -o <path>). Empty
  // disables the rewrite, leaving every op to IREE's own codegen.
  std::string specPath;

  void bindOptions(OptionsBinder &binder) {
    static llvm::cl::OptionCategory category("IREE B200 dispatch kernels");
    binder.opt<std::string>(
        "iree-b200-transform-spec", specPath,
        llvm::cl::desc("Match-and-replace spec binding dense-contraction ops to the "
                       "hand-written sm_100a kernels (hal.dispatch.extern)."),
        llvm::cl::cat(category));
  }
};

struct B200Session : public PluginSession<B200Session, B200Options> {
  void extendPreprocessingPassPipeline(OpPassManager &passManager) override {
    if (options.specPath.empty()) return;
    Preprocessing::InterpreterPassOptions interpreterOptions;
    interpreterOptions.transformSpecPath = options.specPath;
    passManager.addPass(Preprocessing::createInterpreterPass(interpreterOptions));
  }
};

}  // namespace

IREE_DEFINE_COMPILER_OPTION_FLAGS(B200Options);

extern "C" bool iree_register_compiler_plugin_iree_b200(
    mlir::iree_compiler::PluginRegistrar *registrar) {
  registrar->registerPlugin<B200Session>("iree_b200");
  return true;
}
