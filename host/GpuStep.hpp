/*!
 * @file GpuStep.hpp
 * Drop-in replacement for the reference's DevStep (core/src/include/DevStep.hpp:18-36,
 * core/src/DevStep.cpp:14-23) that runs the per-cell column physics on a B200 through the C ABI
 * in include/nsdg_gpu.h.  It implements the reference's own IModelStep interface
 * (core/src/include/IModelStep.hpp:16-34) and is compiled against the reference's own headers;
 * the swap point is the one member declaration `DevStep modelStep;` in
 * core/src/include/Model.hpp:47.
 *
 * State is gathered from the IStructure cursor into structure-of-arrays buffers once
 * (setInitialData), stays resident in HBM across iterate() calls, and is scattered back into
 * the ElementData objects on stop() / writeRestartFile() / sync().
 */
#ifndef HOST_GPUSTEP_HPP
#define HOST_GPUSTEP_HPP

#include "include/IModelStep.hpp"
#include "include/IStructure.hpp"

#include "nsdg_gpu.h"

#include <string>
#include <vector>

namespace Nextsim {

class GpuStep : public IModelStep {
public:
    //! Trivial: the CUDA context is created by setInitialData, so that the reference's
    //! ModuleLoader can hold a static instance next to DevStep on a box without a GPU.
    explicit GpuStep(int device = 0);
    //! Selects the CUDA device; only before setInitialData.
    void setDevice(int device);
    virtual ~GpuStep();
    GpuStep(const GpuStep&) = delete;
    GpuStep& operator=(const GpuStep&) = delete;

    // IModelStep
    void writeRestartFile(const std::string& filePath) override;
    void setInitialData(IStructure& dataStructure) override;
    void init() override {};
    void start(const Iterator::TimePoint& startTime) override {};
    void iterate(const Iterator::Duration& dt) override;
    void stop(const Iterator::TimePoint& stopTime) override;

    //! Re-reads the forcing (ExternalData + wind speed) from the structure and uploads it.
    void updateForcing();
    //! Downloads the device state and writes it into the structure's ElementData.
    void sync();
    //! Number of cells found by the cursor sweep.
    size_t nCells() const { return n; }
    //! The parameter block derived from ModuleLoader + Configurator (for inspection/tests).
    const nsdg_thermo_params& parameters() const { return params; }

private:
    void check(int rc, const char* what) const;
    static nsdg_thermo_params queryConfiguration();

    IStructure* pStructure;
    int device;
    nsdg_ctx* ctx;
    nsdg_thermo_params params;
    size_t n;
    int nLayers;
    bool dirty; // device state is newer than the ElementData
};

} /* namespace Nextsim */

#endif /* HOST_GPUSTEP_HPP */
