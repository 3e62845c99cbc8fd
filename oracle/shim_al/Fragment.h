// TEST INFRASTRUCTURE ONLY: AlleleLikelihoods.h includes Fragment.h but the epilogue path never touches it.
#pragma once
class Fragment {};
