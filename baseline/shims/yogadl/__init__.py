class DataRef: pass
class Storage: pass
class Stream: pass
class Submittable: pass
