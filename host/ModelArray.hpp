/*!
 * @file ModelArray.hpp
 * ModelArray / ModelState / ModelComponent: the field-container trio of later-upstream
 * neXtSIM_DG, which BASELINE.json's north_star names as the interface the GPU path sits behind.
 * NOT IN THE REFERENCE SNAPSHOT (those classes arrive after this
 * snapshot, whose equivalents are ElementData / IStructure / IModelStep).  This header is a
 * minimal, self-contained facade with the upstream names and semantics so that later-upstream
 * modules can bind to the C ABI in include/nsdg_gpu.h:
 *   ModelArray      typed N-d FP64 array with process-wide dimensions per type
 *                   (HField = [ny][nx] cells, VertexField = [ny][nx] CG1 nodes, DG fields with a
 *                   trailing, fastest-varying component dimension as upstream stores them)
 *   ModelState      name -> ModelArray map
 *   ModelComponent  setData(const ModelState&) / getState() / named registration
 * The GPU library keeps DG fields component-major (one coalesced plane per coefficient);
 * toComponentMajor / fromComponentMajor do the transposition at the boundary.
 */
#ifndef HOST_MODELARRAY_HPP
#define HOST_MODELARRAY_HPP

#include <cstddef>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace Nextsim {

class ModelArray {
public:
    enum class Type { H, VERTEX, DG, DGSTRESS, COUNT };
    typedef std::vector<size_t> MultiDim;

    //! Sets the (ny, nx) cell dimensions of every type; VERTEX uses the same periodic topology
    //! (node (i,j) is the lower-left corner of cell (i,j)).
    static void setDimensions(size_t ny, size_t nx)
    {
        dims()[0] = ny;
        dims()[1] = nx;
    }
    //! Number of DG components of a type (H, VERTEX: 1; DG: 1, 3 or 6; DGSTRESS: 3).
    static void setNComponents(Type t, size_t n) { ncomps()[(int)t] = n; }
    static size_t nComponents(Type t) { return ncomps()[(int)t]; }
    static size_t ny() { return dims()[0]; }
    static size_t nx() { return dims()[1]; }

    ModelArray()
        : m_type(Type::H)
    {
    }
    explicit ModelArray(Type t)
        : m_type(t)
        , m_data(ny() * nx() * nComponents(t), 0.)
    {
    }
    static ModelArray HField() { return ModelArray(Type::H); }
    static ModelArray VertexField() { return ModelArray(Type::VERTEX); }
    static ModelArray DGField() { return ModelArray(Type::DG); }
    static ModelArray DGStressField() { return ModelArray(Type::DGSTRESS); }

    Type getType() const { return m_type; }
    size_t nComponents() const { return nComponents(m_type); }
    //! Total number of doubles.
    size_t size() const { return m_data.size(); }
    //! Number of grid points (cells or nodes).
    size_t trueSize() const { return nComponents() ? m_data.size() / nComponents() : 0; }
    MultiDim dimensions() const { return { ny(), nx(), nComponents() }; }

    double& operator[](size_t k) { return m_data[k]; }
    const double& operator[](size_t k) const { return m_data[k]; }
    //! Point (row j, column i), component c (upstream order: component fastest).
    double& operator()(size_t j, size_t i, size_t c = 0) { return m_data[(j * nx() + i) * nComponents() + c]; }
    const double& operator()(size_t j, size_t i, size_t c = 0) const
    {
        return m_data[(j * nx() + i) * nComponents() + c];
    }
    double* data() { return m_data.data(); }
    const double* data() const { return m_data.data(); }

    ModelArray& operator=(double v)
    {
        for (double& x : m_data)
            x = v;
        return *this;
    }

    //! [point][component] (upstream) -> [component][point] (device planes)
    std::vector<double> toComponentMajor() const
    {
        const size_t nc = nComponents(), np = trueSize();
        std::vector<double> out(m_data.size());
        for (size_t p = 0; p < np; ++p)
            for (size_t c = 0; c < nc; ++c)
                out[c * np + p] = m_data[p * nc + c];
        return out;
    }
    void fromComponentMajor(const std::vector<double>& planes)
    {
        const size_t nc = nComponents(), np = trueSize();
        if (planes.size() != m_data.size())
            throw std::length_error("ModelArray::fromComponentMajor: size mismatch");
        for (size_t p = 0; p < np; ++p)
            for (size_t c = 0; c < nc; ++c)
                m_data[p * nc + c] = planes[c * np + p];
    }

private:
    static size_t* dims()
    {
        static size_t d[2] = { 0, 0 };
        return d;
    }
    static size_t* ncomps()
    {
        static size_t n[(int)Type::COUNT] = { 1, 1, 6, 3 };
        return n;
    }
    Type m_type;
    std::vector<double> m_data;
};

typedef ModelArray HField;
typedef std::map<std::string, ModelArray> ModelState;

//! A component of the model that holds or advances named fields.
class ModelComponent {
public:
    virtual ~ModelComponent() { unregisterModule(); }
    virtual std::string getName() const = 0;
    //! Takes the fields it needs out of the state (missing fields keep their current values).
    virtual void setData(const ModelState& state) = 0;
    //! The fields this component owns, current values.
    virtual ModelState getState() const = 0;

    //! Union of the states of all registered components.
    static ModelState getAllModuleStates()
    {
        ModelState all;
        for (auto& kv : registry()) {
            ModelState s = kv.second->getState();
            all.insert(s.begin(), s.end());
        }
        return all;
    }
    static void setAllModuleData(const ModelState& state)
    {
        for (auto& kv : registry())
            kv.second->setData(state);
    }
    static size_t nRegistered() { return registry().size(); }

protected:
    void registerModule() { registry()[getName()] = this; }
    void unregisterModule()
    {
        for (auto it = registry().begin(); it != registry().end();)
            it = (it->second == this) ? registry().erase(it) : ++it;
    }

private:
    static std::map<std::string, ModelComponent*>& registry()
    {
        static std::map<std::string, ModelComponent*> r;
        return r;
    }
};

} /* namespace Nextsim */

#endif /* HOST_MODELARRAY_HPP */
