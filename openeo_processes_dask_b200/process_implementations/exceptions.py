"""Exception types that cross the drop-in boundary.

Same class names and hierarchy as the reference's ``process_implementations/exceptions.py`` (every
error derives from ``OpenEOException``) so ``except`` clauses in host applications keep working.
Only the errors the reducer path can raise are defined; each says where the reference raises it.
"""


class OpenEOException(Exception):
    """Root of all process errors (reference exceptions.py:1)."""


def _error(name: str, doc: str) -> type:
    return type(name, (OpenEOException,), {"__doc__": doc, "__module__": __name__})


ProcessParameterMissing = _error(
    "ProcessParameterMissing",
    "A ParameterReference could not be resolved by the @process decorator (core.py:56-58, 65-67).")
DimensionNotAvailable = _error(
    "DimensionNotAvailable",
    "The requested dimension is not part of the cube (cubes/reduce.py:19-22, cubes/aggregate.py:40-50).")
TooManyDimensions = _error(
    "TooManyDimensions",
    "More than one temporal dimension and none was named (cubes/aggregate.py:51-54).")
QuantilesParameterMissing = _error(
    "QuantilesParameterMissing", "quantiles() got neither `probabilities` nor `q` (math.py:349-352).")
QuantilesParameterConflict = _error(
    "QuantilesParameterConflict", "quantiles() got both `probabilities` and `q` (math.py:344-347).")
DimensionAmbiguous = _error(
    "DimensionAmbiguous", "No (or no unique) dimension of type `bands` (cubes/indices.py:16-19).")
NirBandAmbiguous = _error(
    "NirBandAmbiguous", "The NIR band cannot be resolved by label or common_name (cubes/indices.py:29-36).")
RedBandAmbiguous = _error(
    "RedBandAmbiguous", "The red band cannot be resolved by label or common_name (cubes/indices.py:37-44).")
BandExists = _error(
    "BandExists", "`target_band` names an existing coordinate (cubes/indices.py:54-55).")
ArrayElementParameterMissing = _error(
    "ArrayElementParameterMissing", "array_element() got neither `index` nor `label` (arrays.py:75-78).")
ArrayElementParameterConflict = _error(
    "ArrayElementParameterConflict", "array_element() got both `index` and `label` (arrays.py:80-83).")
ArrayNotLabeled = _error(
    "ArrayNotLabeled", "array_element(label=...) on an array without labels (arrays.py:87-90).")
ArrayElementNotAvailable = _error(
    "ArrayElementNotAvailable", "array_element(): no element with that index or label (arrays.py:117-119).")
DimensionMismatch = _error(
    "DimensionMismatch", "data and mask dimensions do not match (cubes/mask.py:35-37, 63-66).")
DimensionLabelCountMismatch = _error(
    "DimensionLabelCountMismatch", "data and mask have different numbers of temporal labels (cubes/mask.py:46-49).")
LabelMismatch = _error(
    "LabelMismatch", "data and mask temporal labels differ (cubes/mask.py:50-53).")
OverlapResolverMissing = _error(
    "OverlapResolverMissing", "merge_cubes(): the cubes overlap and no overlap resolver was given (cubes/merge.py:169-172, 242-245).")
MinMaxSwapped = _error("MinMaxSwapped", "clip(): min is greater than max (math.py:333-336).")

__all__ = [n for n, v in list(globals().items()) if isinstance(v, type) and issubclass(v, OpenEOException)]
