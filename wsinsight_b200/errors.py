"""Exception names kept from the reference (wsinsight/errors.py:6-47) so callers can catch the same types."""

from __future__ import annotations


class WsinferException(Exception):
    """Base class for wsinsight exceptions."""


class UnknownArchitectureError(WsinferException):
    """Architecture is unknown / the TorchScript graph cannot be lowered to the B200 kernels."""


class WholeSlideImageDirectoryNotFound(WsinferException, FileNotFoundError):
    ...


class DuplicateFilePrefixesFound(WsinferException):
    """Two slides share a stem; stems are the unique slide id."""


class WholeSlideImagesNotFound(WsinferException, FileNotFoundError):
    ...


class ResultsDirectoryNotFound(WsinferException, FileNotFoundError):
    ...


class PatchDirectoryNotFound(WsinferException, FileNotFoundError):
    ...


class CannotReadSpacing(WsinferException):
    ...


class NoBackendException(WsinferException):
    ...


class BackendNotAvailable(WsinferException):
    """The requested backend is not available."""
