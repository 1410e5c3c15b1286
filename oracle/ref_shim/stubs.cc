// Link-time stand-ins for the few CDM / SliceBuilder / CDMReader members that the reference's
// Cached*.cc objects reference from getInputDataSlice()/getOutputDataSlice().  The oracle harness
// never calls those two methods (it drives interpolateValues() directly), so every stub aborts.
#include "fimex/CDM.h"
#include "fimex/CDMReader.h"
#include "fimex/SliceBuilder.h"

#include <cstdio>
#include <cstdlib>

namespace {
[[noreturn]] void unreachable_stub(const char* what)
{
    std::fprintf(stderr, "oracle/_ref stub called: %s\n", what);
    std::abort();
}
} // namespace

namespace MetNoFimex {

SliceBuilder::SliceBuilder(const CDM&, const std::string&, bool) { unreachable_stub("SliceBuilder ctor"); }
SliceBuilder::~SliceBuilder() {}
void SliceBuilder::setStartAndSize(const std::string&, size_t, size_t) { unreachable_stub("SliceBuilder::setStartAndSize"); }
void SliceBuilder::setAll(const std::string&) { unreachable_stub("SliceBuilder::setAll"); }
void SliceBuilder::getStartAndSize(const std::string&, size_t&, size_t&) const { unreachable_stub("SliceBuilder::getStartAndSize"); }
std::vector<std::string> SliceBuilder::getDimensionNames() const { unreachable_stub("SliceBuilder::getDimensionNames"); }
std::vector<std::string> SliceBuilder::getUnsetDimensionNames() const { unreachable_stub("SliceBuilder::getUnsetDimensionNames"); }

const CDMVariable& CDM::getVariable(const std::string&) const { unreachable_stub("CDM::getVariable"); }
const CDMDimension* CDM::getUnlimitedDim() const { unreachable_stub("CDM::getUnlimitedDim"); }
bool CDM::hasUnlimitedDim(const CDMVariable&) const { unreachable_stub("CDM::hasUnlimitedDim"); }

const CDM& CDMReader::getCDM() const { unreachable_stub("CDMReader::getCDM"); }

} // namespace MetNoFimex
