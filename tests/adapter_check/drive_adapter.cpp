#include "b200pixelsim.hpp"
// instantiate everything a caller touches (simulate.cpp:159-181, 89-115, 588-611)
std::size_t drive(const sme::model::Model &m) {
  std::vector<std::string> comps{"c"};
  std::vector<std::vector<std::string>> species{{"A"}};
  std::unique_ptr<sme::simulate::BaseSim> sim{
      std::make_unique<sme::simulate::B200PixelSim>(m, comps, species)};
  std::size_t n{sim->run(1.0, -1.0, [] { return false; })};
  n += sim->getConcentrations(0).size() + sim->getConcentrationPadding();
  sim->setStopRequested(true);
  if (auto *b = dynamic_cast<sme::simulate::B200PixelSim *>(sim.get())) {
    n += b->getDcdt(0).size();
    n += static_cast<std::size_t>(b->getLowerOrderConcentration(0, 0, 0));
  }
  return n + sim->errorMessage().size() + sim->errorImages().volume().nVoxels();
}
