// Host mirror of include/amcl/pf/pdf_gaussian.h of the reference: the RNG contract of the hot path
// (SURVEY.md 8b). draw() is the reference's polar Box-Muller over the process-global POSIX drand48
// (pdf_gaussian.cpp:77-97). sample() needs an eigen-decomposition that the reference takes from
// Eigen::EigenSolver (third party, absent here): a Jacobi decomposition of the symmetric covariance is used
// instead -- same distribution, PARITY UNPINNED against the reference's draws (initialisation only, off the hot path).
#ifndef BAMCL_AMCL_PF_PDF_GAUSSIAN_H
#define BAMCL_AMCL_PF_PDF_GAUSSIAN_H
#include "amcl/compat.h"

namespace badger_amcl
{
class PDFGaussian
{
public:
  PDFGaussian(const Eigen::Vector3d& x, const Eigen::Matrix3d& cx);
  Eigen::Vector3d sample();
  static double draw(double sigma);

  // the decomposition the constructor computed (what bamcl_pf_init_gaussian takes)
  const Eigen::Vector3d& mean() const { return x_; }
  const Eigen::Matrix3d& rotation() const { return cr_; }
  const Eigen::Vector3d& sigmas() const { return cd_; }

private:
  Eigen::Vector3d x_;
  Eigen::Matrix3d cr_;  // rotation (eigenvectors in columns)
  Eigen::Vector3d cd_;  // standard deviations along the eigenvectors
};
}  // namespace badger_amcl
#endif
