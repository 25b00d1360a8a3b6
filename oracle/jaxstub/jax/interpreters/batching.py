class BatchTracer:  # jumpy.py:44
  pass
